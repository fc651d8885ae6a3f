"""Pins the oracle's measurement kernels, lattice iterators and the *sampling* (Metropolis weight) of its sweep
against what the reference's own tests hold for them:

  test/integration_tests.jl:28-93    HubbardModelAttractive(4, 2), beta = 1: <G>, CDC, SDC, PC (K = 5) literals,
                                     atol = 4 dtau^2  (fixtures tests/golden/attractive_4x4_beta1.json)
  test/integration_tests.jl:96-175   HubbardModelRepulsive(2, 2), beta = 1: G, CDC, Mz, SDC, PC (K = 3) literals
                                     (tests/golden/repulsive_2x2_beta1.json)
  test/ED/ED_tests.jl:91-246         2 x 2 exact diagonalisation, atol = rtol = 2 dtau^2 (oracle/ed.py)
  test/lattices.jl / lattice_iterators: counts and ordering of EachSitePairByDistance / EachLocalQuadByDistance
"""
import json
import os

import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from oracle import oracle as orc
from oracle.ed import HubbardED

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _gold(name):
    with open(os.path.join(GOLD, name + ".json")) as f:
        return json.load(f)


# ------------------------------------------------------------------ iterators
@pytest.mark.parametrize("L", [2, 4, 6])
def test_iterator_shapes_and_order(L):
    K = 5 if L > 2 else 3
    t = orc.MeasurementTables(L, K)
    N = L * L
    # every ordered pair exactly once; on a periodic Bravais lattice every direction holds N pairs
    assert len(t.trip) == N * N and {len(p) for p in t.pairs} == {N}
    assert sorted((int(s), int(g)) for _, s, g in t.trip) == [(s, g) for s in range(N) for g in range(N)]
    assert np.allclose(t.directions[0], 0.0)                      # direction 1 = on-site
    for d, lst in enumerate(t.pairs):                             # direction = pos[src] - pos[trg] (wrapped)
        for s, g in lst:
            dx = (s % L - g % L - t.directions[d][0]) % L
            dy = (s // L - g // L - t.directions[d][1]) % L
            assert dx == 0 and dy == 0
    assert all(lst == sorted(lst) for lst in t.pairs)             # origin-major inside a direction
    norms = [np.hypot(*d) for d in t.directions]
    assert all(a <= b + 1e-9 for a, b in zip(norms, norms[1:]))   # sorted by distance
    if L > 2:
        assert [tuple(d) for d in t.directions[1:5]] == [(1, 0), (0, 1), (-1, 0), (0, -1)]  # angle order
    # EachLocalQuadByDistance{K}: length = (sum_src #targets)^2 = (N K)^2 on a Bravais lattice
    assert len(t.quint) == (N * K) ** 2
    assert t.quint[:, 0].max() == t.n_dirs * K * K - 1


@pytest.mark.parametrize("L", [2, 4, 6])
def test_product_tables_agree_with_the_oracle_iterators(L):
    """montecarlo.jl_b200/lattices.py builds the device tables; they must describe the same iterators."""
    t = orc.MeasurementTables(L, 3)
    lat = mcb.SquareLattice(L)
    trg_of, dirs = lat.direction_targets()
    assert trg_of.shape == (t.n_dirs, L * L)
    assert np.allclose(dirs, np.array(t.directions))
    for d, lst in enumerate(t.pairs):
        assert [(s, int(trg_of[d, s])) for s in range(L * L)] == lst


def test_one_flavor_kernels_are_the_two_flavor_kernels_on_a_spin_symmetric_matrix():
    """HubbardModelAttractive.jl:221-245 are the measurements.jl:58-207 kernels for G = blockdiag(g, g)."""
    rng = np.random.default_rng(5)
    t = orc.MeasurementTables(4, 5)
    g = rng.standard_normal((16, 16))
    G2 = np.zeros((32, 32))
    G2[:16, :16] = g
    G2[16:, 16:] = g
    a = orc.measure(g, t, flavors=1)
    b = orc.measure(G2, t, flavors=2)
    for k in ("cdc", "sdc_x", "sdc_y", "sdc_z", "pc", "mx", "my", "mz"):
        assert np.allclose(a[k], b[k], rtol=1e-13, atol=1e-13), k
    assert np.allclose(np.tile(a["occupation"], 2), b["occupation"])
    # the s-wave plane of the pairing correlation (K index 1,1) is sum_i G_{i,i+d}^2 / N
    for d, lst in enumerate(t.pairs):
        assert np.isclose(a["pc"][d, 0, 0], sum(g[s, r] ** 2 for s, r in lst) / 16)


# ------------------------------------------------------------------ statistical goldens of the reference
def _attractive_chains(L, U, mu, beta, dtau, safe_mult, n_chains, seed):
    model = mcb.HubbardModelAttractive(L=L, U=U, mu=mu)
    M = int(round(beta / dtau))
    hop = model.hopping_exponentials(dtau)
    chains = []
    for c in range(n_chains):
        o = orc.OracleDQMC(L * L, M, safe_mult, dtau, U, hop, model.rand_conf(M, seed, c), seed=seed, chain=c)
        o.build_stack()
        chains.append(o)
    return model, chains


def _mean_err(recs, key):
    a = np.stack([np.asarray(r[key], dtype=np.float64) for r in recs])
    return a.mean(0), a.std(0, ddof=1) / np.sqrt(len(recs))


def test_attractive_4x4_beta1_against_the_reference_literals():
    """The reference's run (1 chain x 1000 sweeps, Julia RNG) and ours (independent chains, Philox) sample the same
    distribution: compare means at the reference's own atol = 4 dtau^2, and - much tighter - within the combined
    standard errors (the reference publishes its std_error next to every mean). The SDC rows are the sharp ones:
    their literals carry errors of 3e-5 .. 1.5e-3."""
    g = _gold("attractive_4x4_beta1")
    t = orc.MeasurementTables(4, 5)
    model, chains = _attractive_chains(4, 1.0, 0.0, 1.0, 0.1, 10, n_chains=24, seed=2024)
    recs = [orc.run_measured(o, t, thermalization=10, sweeps=400, measure_rate=10,
                             T=model.hopping_matrix(), U=1.0) for o in chains]
    assert all(r["count"] == 40 for r in recs)
    checks = (("G", "G", lambda x: x.T.reshape(-1)), ("cdc", "CDC", None), ("sdc_x", "SDCx", None),
              ("sdc_y", "SDCy", None), ("sdc_z", "SDCz", None),
              ("pc", "PC", lambda x: x.transpose(2, 1, 0).reshape(-1)))
    for ours, theirs, flat in checks:
        m, e = _mean_err(recs, ours)
        if flat:
            m, e = flat(m), flat(e)
        ref, ref_e = np.array(g[theirs + "_mean"]), np.array(g[theirs + "_std_error"])
        assert m.shape == ref.shape, ours
        # the literal is one noisy sample itself (its G_ii scatter by +-0.05 around the exact 1/2 of half filling)
        assert np.all(np.abs(m - ref) <= g["atol"] + 2 * ref_e), (ours, np.abs(m - ref).max())
        z = np.abs(m - ref) / np.sqrt(e ** 2 + ref_e ** 2 + 1e-8)
        assert z.max() < 6.0 and np.mean(z < 2.5) > 0.9, (ours, z.max(), np.mean(z < 2.5))


def test_repulsive_2x2_beta1_against_the_reference_literals():
    g = _gold("repulsive_2x2_beta1")
    t = orc.MeasurementTables(2, 3)
    model = mcb.HubbardModelRepulsive(L=2, U=1.0)
    hop = model.hopping_exponentials(0.1)
    recs = []
    for c in range(24):
        o = orc.OracleDQMCRepulsive(4, 10, 10, 0.1, 1.0, hop, model.rand_conf(10, 77, c), seed=77, chain=c)
        o.build_stack()
        recs.append(orc.run_measured(o, t, thermalization=10, sweeps=400, measure_rate=10))
    for ours, theirs, flat in (("G", "G", None), ("cdc", "CDC", None), ("mz", "Mz", None), ("sdc_x", "SDCx", None),
                               ("sdc_y", "SDCy", None), ("sdc_z", "SDCz", None),
                               ("pc", "PC", lambda x: x.transpose(2, 1, 0).reshape(-1))):
        m, e = _mean_err(recs, ours)
        if flat:
            m, e = flat(m), flat(e)
        ref, ref_e = np.array(g[theirs + "_mean"]), np.array(g[theirs + "_std_error"])
        assert m.shape == ref.shape, ours
        assert np.abs(m - ref).max() <= g["atol"] + 3 * ref_e.max(), (ours, np.abs(m - ref).max())
        z = np.abs(m - ref) / np.sqrt(e ** 2 + ref_e ** 2 + 1e-8)
        assert z.max() < 6.0, (ours, z.max())


# ------------------------------------------------------------------ exact diagonalisation (test/ED/ED_tests.jl:91-246)
def _close(a, b, tol=0.02):
    return np.all(np.abs(np.asarray(a) - np.asarray(b)) <= tol + tol * np.abs(np.asarray(b)))


def test_attractive_2x2_against_exact_diagonalisation():
    """HubbardModelAttractive(2, 2, U = 1, mu = 1), beta = 1, dtau = 0.1, safe_mult = 5 (ED_tests.jl:93-101)."""
    lat = mcb.SquareLattice(2)
    ed = HubbardED(lat.neighs, U_signed=-1.0, mu=1.0, t=1.0, beta=1.0)
    t = orc.MeasurementTables(2, 4)
    model, chains = _attractive_chains(2, 1.0, 1.0, 1.0, 0.1, 5, n_chains=16, seed=99)
    recs = [orc.run_measured(o, t, thermalization=100, sweeps=1500, measure_rate=5,
                             T=model.hopping_matrix(), U=1.0) for o in chains]
    G_ed = ed.greens_matrix()
    N = 4
    G, _ = _mean_err(recs, "G")
    assert _close(G, G_ed[:N, :N]) and _close(G, G_ed[N:, N:])             # "Greens"
    occ, _ = _mean_err(recs, "occupation")
    assert _close(occ, 1 - np.diag(G_ed)[:N])
    cdc, sdc, pc = np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros((t.n_dirs, 4, 4))
    for d, s, g in t.trip:                                                # ED_tests.jl:158-165, 203-235
        cdc[d] += ed.charge_density_correlation(g, s)
        sdc[d] += ed.spin_density_correlation_z(g, s)
    for lin, s1, t1, s2, t2 in t.quint:                                   # ED_tests.jl:237-246
        pc[lin % t.n_dirs, (lin // t.n_dirs) % 4, lin // (4 * t.n_dirs)] += ed.pairing_correlation(s1, t1, s2, t2)
    assert _close(_mean_err(recs, "cdc")[0], cdc / N)
    for k in ("sdc_x", "sdc_y", "sdc_z"):
        assert _close(_mean_err(recs, k)[0], sdc / N), k
    assert _close(_mean_err(recs, "pc")[0], pc / N)
    # O5: occupation, double occupancy, kinetic and interaction energy per site
    n_ed = np.mean([ed.occupation(i) for i in range(N)])
    d_ed = np.mean([ed.double_occupancy(i) for i in range(N)])
    assert _close(_mean_err(recs, "n")[0], n_ed)
    assert _close(_mean_err(recs, "double_occupancy")[0], d_ed)
    T = model.hopping_matrix()
    ekin_ed = sum(T[i, j] * ed.ev(ed.cd[ed.mode(i, s)] @ ed.c[ed.mode(j, s)])
                  for i in range(N) for j in range(N) for s in (0, 1)) / N
    assert _close(_mean_err(recs, "e_kin")[0], ekin_ed)
    assert _close(_mean_err(recs, "e_int")[0], -1.0 * (d_ed - 0.5 * n_ed + 0.25))


def _ed_current(ed, T, src, trg):
    """current_density (test/ED/ED.jl:403-435): sum over spins of T[trg, src] c+_trg c_src - T[src, trg] c+_src c_trg."""
    J = np.zeros((ed.dim, ed.dim))
    for sub in (0, 1):
        J += T[trg, src] * ed.cd[ed.mode(trg, sub)] @ ed.c[ed.mode(src, sub)]
        J -= T[src, trg] * ed.cd[ed.mode(src, sub)] @ ed.c[ed.mode(trg, sub)]
    return J


def _ed_ccs(ed, T, t, K, dtau):
    """ED_tests.jl:346-358 over EachLocalQuadBySyncedDistance{K}."""
    ccs = np.zeros((t.n_dirs, K))
    for lin, s1, t1, s2, t2 in t.synced:
        ccs[lin % t.n_dirs, lin // t.n_dirs] += ed.integrated(_ed_current(ed, T, s1, t1), _ed_current(ed, T, s2, t2), dtau)
    return ccs


def test_attractive_2x2_susceptibilities_against_exact_diagonalisation():
    """The time-displaced measurements (charge / spin density and pairing susceptibility through the
    CombinedGreensIterator) of the same model against expectation_value_integrated, as ED_tests.jl:272-345 does
    (atol = rtol = 2 dtau^2)."""
    lat = mcb.SquareLattice(2)
    ed = HubbardED(lat.neighs, U_signed=-1.0, mu=1.0, t=1.0, beta=1.0)
    t = orc.MeasurementTables(2, 4)
    N, K, dtau = 4, 4, 0.1
    model, chains = _attractive_chains(2, 1.0, 1.0, 1.0, dtau, 5, n_chains=16, seed=77)
    recs = []
    for o in chains:
        o.propagate_to_measurement_point()
        o.sweep(50)
        acc, n = None, 0
        for _ in range(250):
            o.sweep(2)
            r = orc.susceptibilities(o, t, recalculate=5, T=model.hopping_matrix())
            acc = r if acc is None else {k: acc[k] + r[k] for k in r}
            n += 1
        recs.append({k: v / n for k, v in acc.items()})
    cds, sds, ps = np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros((t.n_dirs, K, K))
    for d, s, g in t.trip:                                                # ED_tests.jl:275-320
        cds[d] += ed.integrated(ed.number(g), ed.number(s), dtau)
        sds[d] += ed.integrated(ed.mz_operator(g), ed.mz_operator(s), dtau)
    for lin, s1, t1, s2, t2 in t.quint:                                   # ED_tests.jl:322-343
        A = ed.c[ed.mode(s1, 0)] @ ed.c[ed.mode(t1, 1)]                   # annihilate (trg1, dn), then (src1, up)
        B = ed.cd[ed.mode(t2, 1)] @ ed.cd[ed.mode(s2, 0)]                 # create (src2, up), then (trg2, dn)
        ps[lin % t.n_dirs, (lin // t.n_dirs) % K, lin // (K * t.n_dirs)] += ed.integrated(A, B, dtau)
    assert _close(_mean_err(recs, "cds")[0], cds / N)
    assert _close(_mean_err(recs, "sds")[0], sds / N)
    assert _close(_mean_err(recs, "ps")[0], ps / N)
    ccs = _ed_ccs(ed, model.hopping_matrix(), t, K, dtau)
    assert _close(_mean_err(recs, "ccs")[0], ccs / N)
    assert np.abs(cds).max() > 0.5 and np.abs(ps).max() > 0.05 and np.abs(ccs).max() > 1.0   # not between zeros


def test_repulsive_2x2_against_exact_diagonalisation():
    """HubbardModelRepulsive(2, 2, U = 1), same run parameters (ED_tests.jl:92)."""
    lat = mcb.SquareLattice(2)
    ed = HubbardED(lat.neighs, U_signed=+1.0, mu=0.0, t=1.0, beta=1.0)
    t = orc.MeasurementTables(2, 4)
    model = mcb.HubbardModelRepulsive(L=2, U=1.0)
    hop = model.hopping_exponentials(0.1)
    recs = []
    for c in range(16):
        o = orc.OracleDQMCRepulsive(4, 10, 5, 0.1, 1.0, hop, model.rand_conf(10, 5, c), seed=5, chain=c)
        o.build_stack()
        recs.append(orc.run_measured(o, t, thermalization=100, sweeps=1500, measure_rate=5))
    G_ed = ed.greens_matrix()
    N = 4
    assert _close(_mean_err(recs, "G")[0], G_ed)
    assert _close(_mean_err(recs, "occupation")[0], 1 - np.diag(G_ed))
    cdc, sdz, sdx, pc = np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros((t.n_dirs, 4, 4))
    for d, s, g in t.trip:
        cdc[d] += ed.charge_density_correlation(g, s)
        sdz[d] += ed.spin_density_correlation_z(g, s)
        sdx[d] += ed.spin_density_correlation_x(g, s)
    for lin, s1, t1, s2, t2 in t.quint:
        pc[lin % t.n_dirs, (lin // t.n_dirs) % 4, lin // (4 * t.n_dirs)] += ed.pairing_correlation(s1, t1, s2, t2)
    assert _close(_mean_err(recs, "cdc")[0], cdc / N)
    assert _close(_mean_err(recs, "sdc_z")[0], sdz / N, 0.03)
    assert _close(_mean_err(recs, "sdc_x")[0], sdx / N, 0.03)
    assert _close(_mean_err(recs, "pc")[0], pc / N)
    mz, mz_e = _mean_err(recs, "mz")
    assert np.all(np.abs(mz - [ed.m_z(i) for i in range(N)]) <= 0.02 + 4 * mz_e)


def test_repulsive_2x2_susceptibilities_against_exact_diagonalisation():
    """The same comparison (ED_tests.jl:272-345) for HubbardModelRepulsive(2, 2, U = 1): the general two-flavor kernels
    for packed Green's functions, including sds_x, which couples the two spin blocks."""
    lat = mcb.SquareLattice(2)
    ed = HubbardED(lat.neighs, U_signed=+1.0, mu=0.0, t=1.0, beta=1.0)
    t = orc.MeasurementTables(2, 4)
    N, K, dtau, M = 4, 4, 0.1, 10
    model = mcb.HubbardModelRepulsive(L=2, U=1.0)
    hop = model.hopping_exponentials(dtau)
    recs = []
    for c in range(16):
        o = orc.OracleDQMCRepulsive(N, M, 5, dtau, 1.0, hop, model.rand_conf(M, 9, c), seed=9, chain=c)
        o.build_stack()
        o.update(M)                      # the measurement point: current_slice == 1 on the way up
        o.sweep(50)
        acc, n = None, 0
        for _ in range(250):
            o.sweep(2)
            r = orc.susceptibilities(o, t, recalculate=5, T=model.hopping_matrix()[:N, :N])
            acc = r if acc is None else {k: acc[k] + r[k] for k in r}
            n += 1
        recs.append({k: v / n for k, v in acc.items()})
    sx_op = lambda s: ed.cd[ed.mode(s, 0)] @ ed.c[ed.mode(s, 1)] + ed.cd[ed.mode(s, 1)] @ ed.c[ed.mode(s, 0)]
    cds, sdx, sdz, ps = np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros((t.n_dirs, K, K))
    for d, s, g in t.trip:
        cds[d] += ed.integrated(ed.number(g), ed.number(s), dtau)
        sdx[d] += ed.integrated(sx_op(g), sx_op(s), dtau)
        sdz[d] += ed.integrated(ed.mz_operator(g), ed.mz_operator(s), dtau)
    for lin, s1, t1, s2, t2 in t.quint:
        A = ed.c[ed.mode(s1, 0)] @ ed.c[ed.mode(t1, 1)]
        B = ed.cd[ed.mode(t2, 1)] @ ed.cd[ed.mode(s2, 0)]
        ps[lin % t.n_dirs, (lin // t.n_dirs) % K, lin // (K * t.n_dirs)] += ed.integrated(A, B, dtau)
    assert _close(_mean_err(recs, "cds")[0], cds / N)
    assert _close(_mean_err(recs, "sds_x")[0], sdx / N, 0.03)
    assert _close(_mean_err(recs, "sds_y")[0], sdx / N, 0.03)
    assert _close(_mean_err(recs, "sds_z")[0], sdz / N, 0.03)
    assert _close(_mean_err(recs, "ps")[0], ps / N)
    ccs = _ed_ccs(ed, model.hopping_matrix()[:N, :N], t, K, dtau)
    assert _close(_mean_err(recs, "ccs")[0], ccs / N)
    assert np.abs(sdx).max() > 0.3 and np.abs(ps).max() > 0.05 and np.abs(ccs).max() > 1.0
