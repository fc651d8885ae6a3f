"""GPU parity of the observables through the C ABI.

  kernel level : same physical G on both sides -> every output of mcb_dqmc_measure against the oracle's restatement of
                 the reference kernels driven by the oracle's own iterator tuple lists, <= 1e-12
  statistics   : HubbardModelAttractive(4, 2), beta = 1 and HubbardModelRepulsive(2, 2) against the literals of the
                 reference's test/integration_tests.jl:28-175 (tests/golden/*.json)
  physics      : 2 x 2 clusters against exact diagonalisation at atol = rtol = 2 dtau^2 (test/ED/ED_tests.jl:91-246)
The oracle side of every one of these is pinned on the CPU in tests/test_oracle_measure.py.
"""
import json
import os

import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from oracle import oracle as orc
from oracle.ed import HubbardED

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _gold(name):
    with open(os.path.join(GOLD, name + ".json")) as f:
        return json.load(f)


def _mean_err(obs, key):
    a = np.asarray(obs[key], dtype=np.float64)
    return a.mean(0), a.std(0, ddof=1) / np.sqrt(a.shape[0])


def _close(a, b, tol=0.02):
    return np.all(np.abs(np.asarray(a) - np.asarray(b)) <= tol + tol * np.abs(np.asarray(b)))


@pytest.mark.parametrize("L,U,beta,n_chains,K", [(2, 1.0, 1.0, 3, 4), (4, 1.0, 1.0, 5, 5), (8, 4.0, 2.0, 4, 5),
                                                 (12, 4.0, 1.0, 3, 5)])
def test_measure_kernel_matches_the_oracle_on_the_same_greens_function(L, U, beta, n_chains, K):
    model = mcb.HubbardModelAttractive(L=L, U=U, mu=0.3)
    mc = mcb.DQMC(model, n_chains=n_chains, seed=17, beta=beta, delta_tau=0.1, safe_mult=5)
    mc.enable_measurements(K)
    mc.build_stack()
    mc.update(7)
    tables = orc.MeasurementTables(L, K)
    T = model.hopping_matrix()
    n_meas = 3
    Gs = []
    for _ in range(n_meas):
        mc.update(3)
        Gs.append(mc.greens(physical=True))
        mc.measure()
    obs = mc.observables()
    assert obs["count"] == n_meas
    for c in range(n_chains):
        want = [orc.measure(G[c], tables, flavors=1, T=T, U=U) for G in Gs]
        avg = lambda k: np.mean([w[k] for w in want], axis=0)
        assert np.abs(obs["G"][c] - np.mean([G[c] for G in Gs], axis=0)).max() < 1e-13
        for k in ("occupation", "cdc", "sdc_x", "sdc_y", "sdc_z", "pc", "n", "double_occupancy", "e_kin", "e_int"):
            assert np.abs(obs[k][c] - avg(k)).max() <= 1e-12 * max(1.0, np.abs(avg(k)).max()), (k, c)
        assert np.allclose(obs["swave_pairing"][c], avg("pc")[:, 0, 0], rtol=0, atol=1e-12)
    # fixed-order sums: a second pass over the same Green's functions is bit-identical
    first = {k: np.array(v) for k, v in obs.items() if k != "count"}
    mc.reset_observables()
    for G in Gs[-1:]:
        mc.measure(); mc.measure()
    a = mc.observables()
    mc.reset_observables()
    mc.measure(); mc.measure()
    b = mc.observables()
    for k in first:
        assert np.array_equal(a[k], b[k]), k
    mc.close()


def test_measure_kernel_two_flavors_matches_the_general_reference_kernels():
    L, U, K = 4, 4.0, 5
    model = mcb.HubbardModelRepulsive(L=L, U=U)
    mc = mcb.DQMC(model, n_chains=3, seed=3, beta=1.0, delta_tau=0.1, safe_mult=5)
    mc.enable_measurements(K)
    mc.build_stack()
    mc.update(9)
    G = mc.greens(physical=True)          # (C, 2, N, N)
    mc.measure()
    obs = mc.observables()
    tables = orc.MeasurementTables(L, K)
    N = L * L
    for c in range(3):
        G2 = np.zeros((2 * N, 2 * N))
        G2[:N, :N], G2[N:, N:] = G[c, 0], G[c, 1]
        want = orc.measure(G2, tables, flavors=2)
        for k in ("occupation", "cdc", "sdc_x", "sdc_y", "sdc_z", "pc", "mz"):
            assert np.abs(obs[k][c] - want[k]).max() <= 1e-12 * max(1.0, np.abs(want[k]).max()), (k, c)
        assert np.abs(obs["G"][c] - G[c]).max() < 1e-14
        nu, nd = 1 - np.diag(G[c, 0]), 1 - np.diag(G[c, 1])
        assert np.isclose(obs["n"][c], (nu + nd).mean(), rtol=1e-13)
        assert np.isclose(obs["double_occupancy"][c], (nu * nd).mean(), rtol=1e-13)
        T = model.hopping_matrix()
        ekin = sum((T * (np.eye(N) - g.T)).sum() for g in G[c]) / N
        assert np.isclose(obs["e_kin"][c], ekin, rtol=1e-12)
        assert np.isclose(obs["e_int"][c], U * ((nu * nd).mean() - 0.5 * (nu + nd).mean() + 0.25), rtol=1e-12)
    mc.close()


def test_run_follows_the_reference_measurement_schedule():
    """run! measures after the update that lands on (slice 1, direction +1) of every measure_rate-th sweep past the
    thermalisation (DQMC.jl:429-436): trajectory parity with the oracle's run loop on the same chains."""
    model = mcb.HubbardModelAttractive(L=4, U=2.0)
    kw = dict(beta=1.0, delta_tau=0.1, safe_mult=5)
    mc = mcb.DQMC(model, n_chains=2, seed=31, thermalization=3, sweeps=9, measure_rate=2, **kw)
    conf0 = mc.conf()
    obs = mc.run()
    tables = orc.MeasurementTables(4, 5)
    for c in range(2):
        o = orc.OracleDQMC(16, 10, 5, 0.1, 2.0, mc.hop, conf0[c], seed=31, chain=c)
        o.build_stack()
        want = orc.run_measured(o, tables, thermalization=3, sweeps=9, measure_rate=2, T=model.hopping_matrix(), U=2.0)
        assert want["count"] == obs["count"] == 5      # sweeps 4, 6, 8, 10, 12
        for k in ("G", "occupation", "cdc", "sdc_x", "pc", "n", "double_occupancy", "e_kin", "e_int"):
            assert np.abs(obs[k][c] - want[k]).max() < 1e-9, (k, c)
    mc.close()


def test_run_schedule_on_the_chain_resident_n144_path():
    """The same on a 12 x 12 lattice: wrap + sweep run inside segment_resident_kernel (runs of up to safe_mult slices, cut at
    the measurement point), the stabilisation through the resident slice chains and the resident blocked QR."""
    model = mcb.HubbardModelAttractive(L=12, U=2.0)
    kw = dict(beta=1.0, delta_tau=0.1, safe_mult=5)
    mc = mcb.DQMC(model, n_chains=2, seed=31, thermalization=1, sweeps=3, measure_rate=1, **kw)
    conf0 = mc.conf()
    obs = mc.run()
    tables = orc.MeasurementTables(12, 5)
    for c in range(2):
        o = orc.OracleDQMC(144, 10, 5, 0.1, 2.0, mc.hop, conf0[c], seed=31, chain=c)
        o.build_stack()
        want = orc.run_measured(o, tables, thermalization=1, sweeps=3, measure_rate=1, T=model.hopping_matrix(), U=2.0)
        assert want["count"] == obs["count"] == 3      # sweeps 2, 3, 4
        for k in ("G", "occupation", "cdc", "sdc_x", "pc", "n", "double_occupancy", "e_kin", "e_int"):
            assert np.abs(obs[k][c] - want[k]).max() < 1e-9, (k, c)
    mc.close()


def test_attractive_4x4_beta1_against_the_reference_literals():
    """test/integration_tests.jl:28-93 with 512 chains x (10 + 1000) sweeps."""
    g = _gold("attractive_4x4_beta1")
    model = mcb.HubbardModelAttractive(L=4, U=1.0)
    mc = mcb.DQMC(model, n_chains=512, seed=123, beta=1.0, delta_tau=0.1, safe_mult=10, thermalization=10,
                  sweeps=1000, measure_rate=10)
    obs = mc.run()
    mc.close()
    assert obs["count"] == 100
    for ours, theirs, flat in (("G", "G", lambda x: x.T.reshape(-1)), ("cdc", "CDC", None), ("sdc_x", "SDCx", None),
                               ("sdc_y", "SDCy", None), ("sdc_z", "SDCz", None),
                               ("pc", "PC", lambda x: x.transpose(2, 1, 0).reshape(-1))):
        m, e = _mean_err(obs, ours)
        if flat:
            m, e = flat(m), flat(e)
        ref, ref_e = np.array(g[theirs + "_mean"]), np.array(g[theirs + "_std_error"])
        assert m.shape == ref.shape
        # the literal is one noisy sample (1 chain, 100 measurements); ours carries errors 20x smaller
        assert np.all(np.abs(m - ref) <= g["atol"] + 2 * ref_e), (ours, np.abs(m - ref).max())
        z = np.abs(m - ref) / np.sqrt(e ** 2 + ref_e ** 2 + 1e-8)
        assert z.max() < 6.0 and np.mean(z < 2.5) > 0.9, (ours, z.max(), np.mean(z < 2.5))
    # exact symmetries of the model at half filling, within our own (small) errors
    m, e = _mean_err(obs, "G")
    assert np.all(np.abs(np.diag(m) - 0.5) < 5 * np.diag(e) + 1e-6)
    m, e = _mean_err(obs, "n")
    assert abs(m - 1.0) < 5 * e + 1e-6


def test_repulsive_2x2_beta1_against_the_reference_literals():
    """test/integration_tests.jl:96-175 (N = 4: the two-sector immediate-update sweep kernel)."""
    g = _gold("repulsive_2x2_beta1")
    model = mcb.HubbardModelRepulsive(L=2, U=1.0)
    mc = mcb.DQMC(model, n_chains=256, seed=123, beta=1.0, delta_tau=0.1, safe_mult=10, thermalization=10,
                  sweeps=1000, measure_rate=10)
    mc.enable_measurements(3)
    obs = mc.run()
    mc.close()
    N = 4
    Gfull = np.zeros((256, 2 * N, 2 * N))
    Gfull[:, :N, :N], Gfull[:, N:, N:] = obs["G"][:, 0], obs["G"][:, 1]
    obs = dict(obs, G=Gfull)
    for ours, theirs, flat in (("G", "G", None), ("cdc", "CDC", None), ("mz", "Mz", None), ("sdc_x", "SDCx", None),
                               ("sdc_y", "SDCy", None), ("sdc_z", "SDCz", None),
                               ("pc", "PC", lambda x: x.transpose(2, 1, 0).reshape(-1))):
        m, e = _mean_err(obs, ours)
        if flat:
            m, e = flat(m), flat(e)
        ref, ref_e = np.array(g[theirs + "_mean"]), np.array(g[theirs + "_std_error"])
        assert m.shape == ref.shape, ours
        assert np.abs(m - ref).max() <= g["atol"] + 3 * ref_e.max(), (ours, np.abs(m - ref).max())
        z = np.abs(m - ref) / np.sqrt(e ** 2 + ref_e ** 2 + 1e-8)
        assert z.max() < 6.0, (ours, z.max())


def test_attractive_2x2_against_exact_diagonalisation():
    """ED_tests.jl:93-246: HubbardModelAttractive(2, 2, U = 1, mu = 1), beta = 1, dtau = 0.1, safe_mult = 5."""
    lat = mcb.SquareLattice(2)
    ed = HubbardED(lat.neighs, U_signed=-1.0, mu=1.0, t=1.0, beta=1.0)
    model = mcb.HubbardModelAttractive(L=2, U=1.0, mu=1.0)
    mc = mcb.DQMC(model, n_chains=256, seed=7, beta=1.0, delta_tau=0.1, safe_mult=5, thermalization=100,
                  sweeps=1000, measure_rate=5)
    mc.enable_measurements(4)
    obs = mc.run()
    mc.close()
    t = orc.MeasurementTables(2, 4)
    N = 4
    G_ed = ed.greens_matrix()
    assert _close(_mean_err(obs, "G")[0], G_ed[:N, :N])
    assert _close(_mean_err(obs, "occupation")[0], 1 - np.diag(G_ed)[:N])
    cdc, sdc, pc = np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros((t.n_dirs, 4, 4))
    for d, s, r in t.trip:
        cdc[d] += ed.charge_density_correlation(r, s)
        sdc[d] += ed.spin_density_correlation_z(r, s)
    for lin, s1, t1, s2, t2 in t.quint:
        pc[lin % t.n_dirs, (lin // t.n_dirs) % 4, lin // (4 * t.n_dirs)] += ed.pairing_correlation(s1, t1, s2, t2)
    assert _close(_mean_err(obs, "cdc")[0], cdc / N)
    assert _close(_mean_err(obs, "sdc_z")[0], sdc / N)
    assert _close(_mean_err(obs, "pc")[0], pc / N)
    n_ed = np.mean([ed.occupation(i) for i in range(N)])
    d_ed = np.mean([ed.double_occupancy(i) for i in range(N)])
    assert _close(_mean_err(obs, "n")[0], n_ed)
    assert _close(_mean_err(obs, "double_occupancy")[0], d_ed)
    T = model.hopping_matrix()
    ekin_ed = sum(T[i, j] * ed.ev(ed.cd[ed.mode(i, s)] @ ed.c[ed.mode(j, s)])
                  for i in range(N) for j in range(N) for s in (0, 1)) / N
    assert _close(_mean_err(obs, "e_kin")[0], ekin_ed)
    assert _close(_mean_err(obs, "e_int")[0], -1.0 * (d_ed - 0.5 * n_ed + 0.25))


def test_repulsive_2x2_against_exact_diagonalisation():
    lat = mcb.SquareLattice(2)
    ed = HubbardED(lat.neighs, U_signed=+1.0, mu=0.0, t=1.0, beta=1.0)
    model = mcb.HubbardModelRepulsive(L=2, U=1.0)
    mc = mcb.DQMC(model, n_chains=256, seed=8, beta=1.0, delta_tau=0.1, safe_mult=5, thermalization=100,
                  sweeps=1000, measure_rate=5)
    mc.enable_measurements(4)
    obs = mc.run()
    mc.close()
    t = orc.MeasurementTables(2, 4)
    N = 4
    G_ed = ed.greens_matrix()
    G, _ = _mean_err(obs, "G")
    assert _close(G[0], G_ed[:N, :N]) and _close(G[1], G_ed[N:, N:])
    assert _close(_mean_err(obs, "occupation")[0], 1 - np.diag(G_ed))
    cdc, sdz, sdx, pc = np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros(t.n_dirs), np.zeros((t.n_dirs, 4, 4))
    for d, s, r in t.trip:
        cdc[d] += ed.charge_density_correlation(r, s)
        sdz[d] += ed.spin_density_correlation_z(r, s)
        sdx[d] += ed.spin_density_correlation_x(r, s)
    for lin, s1, t1, s2, t2 in t.quint:
        pc[lin % t.n_dirs, (lin // t.n_dirs) % 4, lin // (4 * t.n_dirs)] += ed.pairing_correlation(s1, t1, s2, t2)
    assert _close(_mean_err(obs, "cdc")[0], cdc / N)
    assert _close(_mean_err(obs, "sdc_z")[0], sdz / N, 0.03)
    assert _close(_mean_err(obs, "sdc_x")[0], sdx / N, 0.03)
    assert _close(_mean_err(obs, "pc")[0], pc / N)
    mz, mz_e = _mean_err(obs, "mz")
    assert np.all(np.abs(mz - [ed.m_z(i) for i in range(N)]) <= 0.02 + 4 * mz_e)
    d_ed = np.mean([ed.double_occupancy(i) for i in range(N)])
    assert _close(_mean_err(obs, "double_occupancy")[0], d_ed)


@pytest.mark.parametrize("L,U,mu,beta,safe_mult,recalculate,K", [(2, 1.0, 1.0, 1.0, 5, 5, 4), (4, 1.0, 0.3, 2.0, 5, 0, 5),
                                                                   (8, 2.0, 0.0, 2.0, 5, 10, 5), (12, 4.0, 0.0, 1.0, 5, 5, 5)])
def test_susceptibilities_match_the_oracle(L, U, mu, beta, safe_mult, recalculate, K):
    """Time-displaced measurements (charge / spin density and pairing susceptibility, measurements.jl:56-58, 144-146,
    207) accumulated on the device from the device-side CombinedGreensIterator, against the oracle's restatement of
    apply!(::CombinedGreensIterator, ...) on the same HS field; the oracle side is pinned against exact
    diagonalisation in tests/test_oracle_measure.py. Tolerance: the iterator's own (1e-10 on G(0 <- l), G(l <- 0),
    1e-8 on G(l <- l) between recalculations) carried through sums of products of O(1) entries."""
    n_chains = 3
    model = mcb.HubbardModelAttractive(L=L, U=U, mu=mu)
    mc = mcb.DQMC(model, n_chains=n_chains, seed=31, beta=beta, delta_tau=0.1, safe_mult=safe_mult)
    mc.enable_measurements(K)
    conf0 = mc.conf()
    oracles = [orc.OracleDQMC(L * L, mc.M, safe_mult, 0.1, U, mc.hop, conf0[c], seed=31, chain=c) for c in range(n_chains)]
    mc.build_stack()
    with pytest.raises(RuntimeError, match="current_slice == 1"):
        mc.measure_susceptibilities(recalculate)           # not at the measurement point
    mc.update(mc.M)
    for o in oracles:
        o.build_stack()
        o.update(mc.M)
    tables = orc.MeasurementTables(L, K)
    G0, st0 = mc.greens(physical=False), mc.state()
    mc.measure_susceptibilities(recalculate)
    assert np.array_equal(G0, mc.greens(physical=False)) and st0 == mc.state()   # sweep state untouched
    got = mc.susceptibilities()
    assert got["count"] == 1 and got["ps"].shape == (n_chains, L * L, K, K)
    T = model.hopping_matrix()
    want = [orc.susceptibilities(o, tables, recalculate, T=T) for o in oracles]
    assert got["ccs"].shape == (n_chains, L * L, K)
    for c in range(n_chains):
        for k in ("cds", "sds", "ps", "ccs"):
            scale = max(1.0, np.abs(want[c][k]).max())
            assert np.abs(got[k][c] - want[c][k]).max() < 1e-8 * scale, (c, k)
    assert np.abs(want[0]["cds"]).max() > 0.1 and np.abs(want[0]["ccs"]).max() > 0.1
    # a second measurement after two more sweeps: the record is the mean over calls
    mc.sweep(2)
    mc.measure_susceptibilities(recalculate)
    two = mc.susceptibilities()
    assert two["count"] == 2
    for c, o in enumerate(oracles):
        o.sweep(2)
        w2 = orc.susceptibilities(o, tables, recalculate, T=T)
        for k in ("cds", "sds", "ps", "ccs"):
            scale = max(1.0, np.abs(w2[k]).max())
            assert np.abs(two[k][c] - 0.5 * (want[c][k] + w2[k])).max() < 1e-8 * scale, (c, k)
    mc.reset_observables()
    assert mc.susceptibilities()["count"] == 0
    mc.close()


def test_scheduled_susceptibilities_ride_along_with_the_measurement_schedule():
    """run!'s measurement point (DQMC.jl:423-438): with schedule_susceptibilities the time-displaced group is measured
    wherever the equal-time one is; the sums equal those of the same measurements taken by hand."""
    def make():
        mc = mcb.DQMC(mcb.HubbardModelAttractive(L=4, U=2.0, mu=0.2), n_chains=3, seed=5, beta=1.0, delta_tau=0.1, safe_mult=5)
        mc.enable_measurements(5)
        mc.build_stack()
        mc.update(mc.M)           # to the measurement point; sweeps counted from here end there
        return mc
    a = make()
    a.schedule_susceptibilities(True, recalculate=5)
    a.sweep(6, measure_every=2)
    sa, oa = a.susceptibilities(), a.observables()
    b = make()
    for k in range(1, 7):
        b.sweep(1)
        # the update that ends a visit of the measurement point belongs to sweep k + 1 here (the handle started half a
        # sweep in): DQMC.jl:420-423 counts sweeps from the start of the run
        if (k + 1) % 2 == 0:
            b.measure()
            b.measure_susceptibilities(recalculate=5)
    sb, ob = b.susceptibilities(), b.observables()
    assert sa["count"] == sb["count"] == oa["count"] == 3
    for k in ("cds", "sds", "ps", "ccs"):
        assert np.array_equal(sa[k], sb[k]), k
    assert np.array_equal(oa["cdc"], ob["cdc"])
    a.close()
    b.close()


def test_replay_takes_the_time_displaced_group_along():
    """replay! (DQMC.jl:649-660) with a measurement group on the CombinedGreensIterator: per recorded configuration
    calculate_greens(mc, 0), then the iterator from that G(0). Reference values: the same kernels on the STABILISED
    G(0 <- l), G(l <- 0), G(l <- l) of every slice (greens(mc, k, l)), which the iterator reproduces to 1e-10 / 1e-8."""
    L, U, K = 4, 2.0, 5
    model = mcb.HubbardModelAttractive(L=L, U=U, mu=0.1)
    kw = dict(beta=1.0, delta_tau=0.1, safe_mult=5)
    mc = mcb.DQMC(model, n_chains=3, seed=9, **kw)
    mc.enable_measurements(K)
    mc.schedule_susceptibilities(True, recalculate=5)
    N, M = mc.N, mc.M
    confs = np.random.default_rng(4).choice(np.array([-1, 1], dtype=np.int8), size=(6, N, M))
    res = mc.replay(np.stack([mcb.compress(c) for c in confs]))
    assert res[0]["count"] == 2
    got = mc.susceptibilities()
    assert got["count"] == 2
    tables = orc.MeasurementTables(L, K)
    one = mcb.DQMC(model, n_chains=1, seed=9, **kw)
    want = []
    for c in confs:
        one.set_conf(c[None])
        one.greens_at(0)                                     # leaves G(0) as the handle's Green's function
        G00 = one.greens(physical=True)[0]
        cg = np.stack([np.stack([one.greens_unequal(0, l)[0], one.greens_unequal(l, 0)[0], one.greens_unequal(l, l)[0]])
                       for l in range(1, M + 1)])
        want.append(orc.susceptibilities((G00, cg, 0.1), tables, T=model.hopping_matrix()))
    for slot in range(3):
        for k in ("cds", "sds", "ps", "ccs"):
            w = 0.5 * (want[slot][k] + want[3 + slot][k])    # batch b holds configurations 3 b .. 3 b + 2
            assert np.abs(got[k][slot] - w).max() < 1e-7 * max(1.0, np.abs(w).max()), (slot, k)
    mc.close(); one.close()
