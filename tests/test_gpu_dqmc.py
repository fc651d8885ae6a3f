"""GPU parity tests of the DQMC path: libmcb200 (through the C ABI) against the CPU oracle.

P1 kernel level  : vmul / udt / rdivp / calculate_greens on random inputs
                   (test/slice_matrices.jl:141-235 restated on the batched entry points)
P2 state level   : same (hopping, conf, Philox stream) -> after every update() of a sweep
                   max|G_gpu - G_oracle| / max|G_oracle| <= 1e-10 and identical HS fields
                   (tolerance from BASELINE.json north_star)
"""
import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import _lib
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
RTOL_G = 1e-10  # north_star: per-slice Green's functions within 1e-10 relative max-norm


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def _cm(a):  # batch of numpy matrices -> [batch][col][row] buffer
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).transpose(0, 2, 1)).reshape(-1)


def _uncm(buf, b, n):
    return buf.reshape(b, n, n).transpose(0, 2, 1).copy()


@pytest.mark.parametrize("n", [4, 16, 37, 64, 144])
def test_vmul_modes(n):
    rng = np.random.default_rng(n)
    b = 3
    A, B = rng.standard_normal((b, n, n)), rng.standard_normal((b, n, n))
    L = _lib.lib()
    for mode, want in ((0, A @ B), (1, A @ B.transpose(0, 2, 1)), (2, A.transpose(0, 2, 1) @ B)):
        out = np.empty(b * n * n)
        _lib.check(L.mcb_linalg_vmul(n, b, _cm(A), _cm(B), out, mode, 0))
        assert relerr(_uncm(out, b, n), want) < 1e-13
        for k in range(b):
            assert relerr(_uncm(out, b, n)[k], orc.vmul(A[k], B[k], mode)) < 1e-13


def test_vmul_whole_matrix_kernel_n144():
    """Batches that fill the device at N = 144 take the one-CTA-per-matrix GEMM (gemm_full_kernel): all transpose
    modes against numpy and the oracle's loop nests."""
    n, b = 144, 96
    rng = np.random.default_rng(5)
    A, B = rng.standard_normal((b, n, n)), rng.standard_normal((b, n, n))
    L = _lib.lib()
    for mode, want in ((0, A @ B), (1, A @ B.transpose(0, 2, 1)), (2, A.transpose(0, 2, 1) @ B)):
        out = np.empty(b * n * n)
        _lib.check(L.mcb_linalg_vmul(n, b, _cm(A), _cm(B), out, mode, 0))
        assert relerr(_uncm(out, b, n), want) < 1e-13
        assert relerr(_uncm(out, b, n)[7], orc.vmul(A[7], B[7], mode)) < 1e-13


@pytest.fixture
def linalg_path(request):
    """general=True forces the any-N kernels where a chain-resident kernel exists (N = 16, 32, 64)."""
    L = _lib.lib()
    L.mcb_linalg_force_general(int(request.param))
    yield request.param
    L.mcb_linalg_force_general(0)


@pytest.mark.parametrize("linalg_path", [False, True], indirect=True)
@pytest.mark.parametrize("n", [4, 16, 32, 64, 100, 144, 200, 256, 400])
@pytest.mark.parametrize("apply_pivot", [1, 0])
def test_udt_contract(n, apply_pivot, linalg_path):
    rng = np.random.default_rng(100 + n)
    b = 2
    X = rng.standard_normal((b, n, n))
    X[1] = X[1] @ np.diag(np.logspace(0, -12, n))  # graded columns, like a long B-chain
    L = _lib.lib()
    U, T = np.empty(b * n * n), np.empty(b * n * n)
    D, piv = np.empty(b * n), np.empty(b * n, dtype=np.int32)
    _lib.check(L.mcb_linalg_udt(n, b, _cm(X), U, D, T, piv, apply_pivot, 0))
    U, T = _uncm(U, b, n), _uncm(T, b, n)
    D, piv = D.reshape(b, n), piv.reshape(b, n)
    for k in range(b):
        assert np.abs(U[k] @ U[k].T - np.eye(n)).max() < 1e-12
        scale = np.abs(X[k]).max()
        if apply_pivot:
            assert np.abs(U[k] @ np.diag(D[k]) @ T[k] - X[k]).max() < 1e-12 * scale * n
        else:
            P = np.zeros((n, n))
            P[np.arange(n), piv[k]] = 1.0
            assert np.abs(U[k] @ np.diag(D[k]) @ np.triu(T[k]) @ P - X[k]).max() < 1e-12 * scale * n
        Uo, Do, To, po = orc.udt_pivot(X[k], bool(apply_pivot))
        assert sorted(piv[k]) == list(range(n))
        if np.array_equal(po, piv[k]):
            assert np.allclose(D[k], Do, rtol=1e-9)
            assert np.abs(U[k] - Uo).max() < 1e-9
        else:
            # panel-pivoted kernels (n % 16 == 0, n >= 128, csrc/udt_blocked.cu) choose 16 columns at a time:
            # the factorisation is a different one, with the same determinant and the same scale separation
            assert not linalg_path and n % 16 == 0 and n >= 128
            assert np.isclose(np.log(D[k]).sum(), np.log(Do).sum(), rtol=1e-9, atol=1e-9)
            Tt = np.triu(T[k]) if not apply_pivot else T[k]
            assert np.abs(Tt).max() < 8.0                      # strict pivoting bounds |T_ij| by 1
            assert np.linalg.cond(Tt) < 50 * np.linalg.cond(np.triu(To) if not apply_pivot else To)
            lead = D[k].reshape(-1, 16)[:, 0]                  # the largest column of every panel: non-increasing
            assert np.all(np.diff(lead) <= 1e-9 * lead[0])
            # the first panel = the 16 columns of largest norm, in decreasing order
            assert np.array_equal(piv[k][:16], np.argsort(-(X[k] ** 2).sum(0), kind="stable")[:16])


@pytest.mark.parametrize("n", [144, 256, 400])
def test_udt_equal_norms(n):
    """Pivot selection when every column norm is EXACTLY equal (identity, a +-1 matrix): every column passes the
    selection's threshold, ties go to the lower position, the factorisation still reconstructs."""
    rng = np.random.default_rng(n)
    b = 2
    X = np.empty((b, n, n))
    X[0] = np.eye(n)
    X[1] = rng.choice([-1.0, 1.0], size=(n, n))
    L = _lib.lib()
    U, T = np.empty(b * n * n), np.empty(b * n * n)
    D, piv = np.empty(b * n), np.empty(b * n, dtype=np.int32)
    _lib.check(L.mcb_linalg_udt(n, b, _cm(X), U, D, T, piv, 1, 0))
    U, T = _uncm(U, b, n), _uncm(T, b, n)
    D, piv = D.reshape(b, n), piv.reshape(b, n)
    for k in range(b):
        assert sorted(piv[k]) == list(range(n))
        assert np.array_equal(piv[k][:16], np.arange(16))
        assert np.abs(U[k] @ U[k].T - np.eye(n)).max() < 1e-12
        assert np.abs(U[k] @ np.diag(D[k]) @ T[k] - X[k]).max() < 1e-12 * n
    assert np.array_equal(piv[0], np.arange(n))   # the identity never moves a column


@pytest.mark.parametrize("linalg_path", [False, True], indirect=True)
@pytest.mark.parametrize("n", [4, 16, 64, 144, 272, 400])   # 272, 400: blocked substitution on the tensor pipe
def test_rdivp(n, linalg_path):
    rng = np.random.default_rng(7 + n)
    b = 2
    X = rng.standard_normal((b, n, n))
    A = rng.standard_normal((b, n, n))
    L = _lib.lib()
    out = _cm(A).copy()
    Ts, Ps = [], []
    for k in range(b):
        _, _, T, p = orc.udt_pivot(X[k], False)
        Ts.append(T)
        Ps.append(p)
    _lib.check(L.mcb_linalg_rdivp(n, b, out, _cm(np.array(Ts)), np.array(Ps, dtype=np.int32).reshape(-1), 0))
    got = _uncm(out, b, n)
    for k in range(b):
        assert relerr(got[k], orc.rdivp(A[k], Ts[k], Ps[k])) < 1e-10


@pytest.mark.parametrize("linalg_path", [False, True], indirect=True)
@pytest.mark.parametrize("n", [16, 32, 64])
def test_calculate_greens_from_udts(n, linalg_path):
    rng = np.random.default_rng(n)
    b = 2
    trip = []
    for side in range(2):
        Us, Ds, Ts = [], [], []
        for k in range(b):
            X = rng.standard_normal((n, n)) @ np.diag(np.logspace(2, -6, n))
            U, D, T, _ = orc.udt_pivot(X, True)
            Us.append(U); Ds.append(D); Ts.append(T)
        trip.append((np.array(Us), np.array(Ds), np.array(Ts)))
    (Ul, Dl, Tl), (Ur, Dr, Tr) = trip
    L = _lib.lib()
    G = np.empty(b * n * n)
    _lib.check(L.mcb_linalg_calculate_greens(n, b, _cm(Ul), np.ascontiguousarray(Dl).reshape(-1), _cm(Tl),
                                             _cm(Ur), np.ascontiguousarray(Dr).reshape(-1), _cm(Tr), G, 0))
    G = _uncm(G, b, n)
    for k in range(b):
        want = orc.calculate_greens_udt(Ul[k], Dl[k], Tl[k], Ur[k], Dr[k], Tr[k])
        assert relerr(G[k], want) < RTOL_G
        direct = np.linalg.inv(np.eye(n) + Ul[k] @ np.diag(Dl[k]) @ Tl[k] @ Tr[k].T @ np.diag(Dr[k]) @ Ur[k].T)
        assert relerr(G[k], direct) < 1e-8


def _pair(L, U, beta, dtau, safe_mult, n_chains, seed=123, check=False, chain_offset=0, general=False):
    model = mcb.HubbardModelAttractive(L=L, U=U)
    mc = mcb.DQMC(model, n_chains=n_chains, seed=seed, chain_offset=chain_offset, beta=beta, delta_tau=dtau,
                  safe_mult=safe_mult, check_propagation_error=check, general_path=general)
    conf0 = mc.conf()
    oracles = [orc.OracleDQMC(L * L, mc.M, safe_mult, dtau, U, mc.hop, conf0[c], seed=seed,
                              chain=chain_offset + c, check_propagation=check) for c in range(n_chains)]
    return mc, oracles


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("L,U,beta,safe_mult,chains", [(2, 1.0, 1.0, 5, 2), (4, 1.0, 1.0, 5, 3), (4, 4.0, 2.0, 10, 2)])
def test_every_update_of_a_sweep_matches_oracle(L, U, beta, safe_mult, chains, general):
    mc, oracles = _pair(L, U, beta, 0.1, safe_mult, chains, check=True, general=general)
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    G = mc.greens(physical=False)
    for c, o in enumerate(oracles):
        assert relerr(G[c], o.greens()) < RTOL_G
    for u in range(2 * mc.M + 3):
        mc.update(1)
        G = mc.greens(physical=False)
        st = mc.state()
        conf = mc.conf()
        for c, o in enumerate(oracles):
            o.update(1)
            so = o.state()
            assert (st["current_slice"], st["direction"]) == (so["current_slice"], so["direction"])
            assert relerr(G[c], o.greens()) < RTOL_G, (u, c)
            assert np.array_equal(conf[c], o.conf()), (u, c)
    stats = mc.stats()
    for c, o in enumerate(oracles):
        so = o.state()
        assert stats["acc_local"][c] == so["acc_local"] and stats["prop_local"][c] == so["prop_local"]
        assert stats["max_propagation_error"][c] < 1e-7
    Gp = mc.greens(physical=True)
    for c, o in enumerate(oracles):
        assert relerr(Gp[c], o.greens(physical=True)) < RTOL_G
    mc.close()


@pytest.mark.parametrize("general", [False, True])
def test_config_a_8x8_beta8_full_sweep_matches_oracle(general):
    """BASELINE.json configs[0]/[1] physics: 8x8, U=1, beta=8, dtau=0.1, safe_mult=10."""
    mc, oracles = _pair(8, 1.0, 8.0, 0.1, 10, 2, general=general)
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    checkpoints = [1, 9, 10, 11, 79, 80, 81, 120, 159, 160]
    done = 0
    for cp in checkpoints:
        mc.update(cp - done)
        G = mc.greens(physical=False)
        conf = mc.conf()
        for c, o in enumerate(oracles):
            o.update(cp - done)
            assert relerr(G[c], o.greens()) < RTOL_G, (cp, c)
            assert np.array_equal(conf[c], o.conf()), (cp, c)
        done = cp
    mc.close()


def test_config_c_12x12_u4_beta10_partial_sweep_matches_oracle():
    """BASELINE.json configs[2] physics (N=144): first 25 updates incl. two stabilisations."""
    mc, oracles = _pair(12, 4.0, 10.0, 0.1, 10, 2)
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    done = 0
    for cp in (1, 10, 11, 25):
        mc.update(cp - done)
        G = mc.greens(physical=False)
        conf = mc.conf()
        for c, o in enumerate(oracles):
            o.update(cp - done)
            assert relerr(G[c], o.greens()) < RTOL_G, (cp, c)
            assert np.array_equal(conf[c], o.conf()), (cp, c)
        done = cp
    mc.close()


def test_greens_from_scratch_at_every_slice():
    """test/flavortests_DQMC.jl:60-69: calculate_greens(mc, k) for k in 0:M against the oracle."""
    mc, oracles = _pair(4, 1.0, 2.0, 0.1, 5, 2)
    for k in range(0, mc.M + 1):
        G = mc.greens_at(k)
        for c, o in enumerate(oracles):
            assert relerr(G[c], o.greens_at(k)) < RTOL_G, k
    mc.close()


@pytest.mark.parametrize("L", [4, 8])
def test_batched_updates_equal_single_updates(L):
    """update(n) runs many slice visits per launch on the chain-resident path; the trajectory must be
    bit-identical to n single update() calls and (HS field) to the general any-N kernels."""
    a, _ = _pair(L, 2.0, 2.0, 0.1, 5, 3, seed=5)
    b, _ = _pair(L, 2.0, 2.0, 0.1, 5, 3, seed=5)
    c, _ = _pair(L, 2.0, 2.0, 0.1, 5, 3, seed=5, general=True)
    for mc in (a, b, c):
        mc.build_stack()
    n = 2 * a.M + 7
    a.update(n)
    for _ in range(n):
        b.update(1)
    c.update(n)
    assert a.state() == b.state() == c.state()
    assert np.array_equal(a.conf(), b.conf()) and np.array_equal(a.greens(False), b.greens(False))
    assert np.array_equal(a.conf(), c.conf())
    assert relerr(a.greens(False), c.greens(False)) < RTOL_G
    sa, sc = a.stats(), c.stats()
    assert np.array_equal(sa["acc_local"], sc["acc_local"]) and np.array_equal(sa["prop_local"], sc["prop_local"])
    for mc in (a, b, c):
        mc.close()


def test_chain_ids_are_global():
    """Chains keyed by global id: a handle with chain_offset=2 reproduces chains 2,3 of a 4-chain handle."""
    full, _ = _pair(4, 1.0, 1.0, 0.1, 5, 4)
    part, _ = _pair(4, 1.0, 1.0, 0.1, 5, 2, chain_offset=2)
    assert np.array_equal(full.conf()[2:], part.conf())
    for mc in (full, part):
        mc.build_stack()
        mc.sweep(1)
    assert np.array_equal(full.conf()[2:], part.conf())
    assert np.array_equal(full.greens()[2:], part.greens())
    full.close(); part.close()


def test_errors_are_reported():
    L = _lib.lib()
    model = mcb.HubbardModelAttractive(L=4)
    with pytest.raises(ValueError):
        mcb.DQMC(model, beta=1.0, delta_tau=0.1, safe_mult=3)
    mc = mcb.DQMC(model, beta=1.0, delta_tau=0.1, safe_mult=5)
    with pytest.raises(mcb.McbError):
        mc.update(1)  # build_stack first
    mc.close()


def test_sharded_run_equals_plain_run_on_one_gpu():
    """ShardedDQMC (the multi-GPU driver) with world_size 1: same chains, same observables."""
    model = mcb.HubbardModelAttractive(L=4, U=2.0)
    kw = dict(beta=1.0, delta_tau=0.1, safe_mult=5, thermalization=2, sweeps=6, measure_rate=2)
    mc = mcb.DQMC(model, n_chains=6, seed=5, **kw)
    obs = mc.run()
    mc.close()
    sh = mcb.ShardedDQMC(model, 6, seed=5, **kw)
    res, n = sh.run()
    sh.close()
    assert n == 6
    for key in ("occupation", "double_occupancy", "energy", "cdc", "swave_pairing", "G"):
        mean, err = mcb.mean_and_error(obs[key])
        assert np.allclose(res[key][0], mean, rtol=1e-12, atol=1e-14)
        assert np.allclose(res[key][1], err, rtol=1e-8, atol=1e-14)


def test_full_size_config_a_512_chains_properties():
    """BASELINE.json configs[1] at full size (512 chains): size-independent properties.
    (1) the propagated G agrees with the from-scratch G of the same HS field (the reference's own
    canary, stack.jl:540-550) for every chain; (2) chains c of the big batch reproduce the oracle run
    alone; (3) no chain produced a non-finite G; (4) the batch equals two half batches."""
    model = mcb.HubbardModelAttractive(L=8, U=1.0)
    kw = dict(beta=8.0, delta_tau=0.1, safe_mult=10, check_propagation_error=True)
    mc = mcb.DQMC(model, n_chains=512, seed=123, **kw)
    conf0 = mc.conf()
    mc.build_stack()
    mc.sweep(1)
    st = mc.stats()
    assert st["max_propagation_error"].max() < 1e-7
    assert st["n_nonfinite"].sum() == 0
    assert (st["prop_local"] == 2 * mc.M * mc.N).all()
    G = mc.greens(physical=False)
    conf = mc.conf()
    state = mc.state()
    assert (state["current_slice"], state["direction"]) == (mc.M, -1)
    fresh = mc.greens_at(mc.M - 1)  # G_M = [1 + B_{M-1} ... B_1 B_M]^-1 from scratch
    assert relerr(G, fresh) < 1e-8
    for c in (0, 255, 511):
        o = orc.OracleDQMC(64, mc.M, 10, 0.1, 1.0, mc.hop, conf0[c], seed=123, chain=c)
        o.build_stack()
        o.update(2 * mc.M)
        assert np.array_equal(conf[c], o.conf())
        assert relerr(G[c], o.greens()) < RTOL_G
    mc.close()
    half = mcb.DQMC(model, n_chains=256, seed=123, chain_offset=256, **kw)
    half.build_stack()
    half.sweep(1)
    assert np.array_equal(half.conf(), conf[256:])
    assert np.array_equal(half.greens(physical=False), G[256:])
    half.close()


def test_config_d_20x20_first_updates_match_oracle():
    """BASELINE.json configs[3] physics (N=400, U=4, dtau=0.05, safe_mult=10) at a reduced number of
    slices so that the CPU oracle finishes in seconds: the general any-N kernels incl. the row-block
    rdivp and the delayed-update sweep with K < 32."""
    mc, oracles = _pair(20, 4.0, 1.0, 0.05, 10, 2)
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    done = 0
    for cp in (1, 10, 11):
        mc.update(cp - done)
        G = mc.greens(physical=False)
        conf = mc.conf()
        for c, o in enumerate(oracles):
            o.update(cp - done)
            assert relerr(G[c], o.greens()) < RTOL_G, (cp, c)
            assert np.array_equal(conf[c], o.conf()), (cp, c)
        done = cp
    mc.close()


def test_blockwise_sweep_scheme_on_small_lattices():
    """The block-wise decision scheme of the delayed sweep (decide_block + triangular recurrences, sweep_delayed.cu)
    is the default for N > 64 (config C and D tests above). MCB_SWEEP_EXP=32 forces it for every size: the
    every-update parity tests of the general path (4x4, 8x8 full sweep) and config C must hold with it too. The switch
    is read once per process, hence the child process."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, MCB_SWEEP_EXP="32")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_dqmc.py"), "-m", "gpu", "-q",
                        "-x", "-k", "(every_update and True) or (config_a and True) or config_c"],
                       env=env, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "no tests ran" not in r.stdout


# ------------------------------------------------------------------ unequal-time Green's functions
UNEQUAL_PAIRS = [(0, 0), (3, 3), (7, 2), (6, 5), (10, 4), (2, 7), (4, 5), (0, 3), (5, 10)]


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("L,U,beta,safe_mult", [(4, 1.0, 2.0, 5), (8, 2.0, 2.0, 10), (6, 4.0, 2.0, 4)])
def test_unequal_time_greens_matches_oracle(L, U, beta, safe_mult, general):
    """greens(mc, k, l) (unequal_time_stack.jl:291-305, full1 / full2) against the CPU restatement, on the
    chain-resident kernels (N = 16, 64), the general kernels, and a size without panel kernels (N = 36)."""
    mc, oracles = _pair(L, U, beta, 0.1, safe_mult, 2, general=general)
    M = mc.M
    for k, l in UNEQUAL_PAIRS + [(M, 0), (0, M), (M, M), (M - 1, 1), (1, M - 1)]:
        G = mc.greens_unequal(k, l, physical=False)
        for c, o in enumerate(oracles):
            want = o.greens_unequal(k, l)
            assert relerr(G[c], want) < RTOL_G, (k, l, c, relerr(G[c], want))
    Gp = mc.greens_unequal(7, 2, physical=True)
    assert relerr(Gp[0], oracles[0].greens_unequal(7, 2, physical=True)) < RTOL_G
    mc.close()


def test_unequal_time_properties_and_sweep_state():
    """test/flavortests_DQMC.jl:103-115 at beta = 15, safe_mult = 5: G(k, k) equals calculate_greens(mc, k),
    G(t, 0) = -G(t, beta); the call leaves the sweep state alone and follows the configuration after updates."""
    mc, oracles = _pair(4, 1.0, 15.0, 0.1, 5, 2, seed=11)
    M = mc.M
    for s in (0, 1, 5, 6, 75, 149, 150):
        assert np.abs(mc.greens_at(s) - mc.greens_unequal(s, s, physical=False)).max() < 1e-11, s
    for s in (0, 1, 37, 100, 149):
        assert np.abs(mc.greens_unequal(s, 0, physical=False) + mc.greens_unequal(s, M, physical=False)).max() < 1e-11, s
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    mc.update(7)
    G0, st0 = mc.greens(physical=False), mc.state()
    Gu = mc.greens_unequal(20, 4, physical=False)
    assert np.array_equal(G0, mc.greens(physical=False)) and st0 == mc.state()
    for c, o in enumerate(oracles):
        o.update(7)
        assert relerr(Gu[c], o.greens_unequal(20, 4)) < RTOL_G  # stacks rebuilt for the updated field
    mc.update(3)  # the sweep carries on
    assert mc.state()["n_updates"] == st0["n_updates"] + 3
    mc.close()


def _to_measurement_point(mc, oracles):
    """build_stack + M updates: current_slice == 1 on the way up, mc.s.greens = G(0) (where DQMC.jl's run loop measures)."""
    mc.build_stack()
    mc.update(mc.M)
    for o in oracles:
        o.build_stack()
        o.update(mc.M)
    st = mc.state()
    assert st["current_slice"] == 1 and st["direction"] == 1


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("L,U,beta,safe_mult,recalculate", [(4, 1.0, 2.0, 5, 0), (4, 1.0, 2.0, 5, 5), (8, 2.0, 4.0, 4, 0),
                                                            (8, 4.0, 2.0, 5, 10), (6, 4.0, 2.4, 4, 8), (2, 1.0, 15.0, 5, 0)])
def test_combined_greens_iterator(L, U, beta, safe_mult, recalculate, general):
    """CombinedGreensIterator on the device (unequal_time_stack.jl:746-885) against its step-for-step CPU restatement
    and, as test/flavortests_DQMC.jl:143-160 does, against greens(mc, 0, l), greens(mc, l, 0), greens(mc, l, l).
    The configurations are ones where the reference's scheme itself is accurate: at 8 x 8, beta = 4 with safe_mult = 10
    and the default recalculate = 40 = M the restated reference is off by 5e-7 in G(0 <- l) and by O(10) in G(l <- l)
    at l = 39 (39 unstabilised similarity transforms)"""
    mc, oracles = _pair(L, U, beta, 0.1, safe_mult, 2, general=general)
    _to_measurement_point(mc, oracles)
    G0, st0 = mc.greens(physical=False), mc.state()
    it = mcb.CombinedGreensIterator(mc, recalculate=recalculate)
    assert len(it) == mc.M
    got = np.stack([np.stack(t, axis=1) for t in it])           # (M, C, 3, N, N)
    assert got.shape == (mc.M, 2, 3, L * L, L * L)
    assert np.array_equal(G0, mc.greens(physical=False)) and st0 == mc.state()   # sweep state untouched
    direct = np.stack([np.stack([mc.greens_unequal(0, l), mc.greens_unequal(l, 0), mc.greens_unequal(l, l)], axis=1)
                       for l in range(1, mc.M + 1)])
    # recalculate = safe_mult: every stabilisation is a recalculation ("high precision"); otherwise G(l <- l) between
    # recalculations is a similarity transform that amplifies rounding (see tests/test_oracle_unequal.py)
    tol_ll = 1e-11 if recalculate == safe_mult else 1e-8
    assert np.abs(got[:, :, :2] - direct[:, :, :2]).max() < 1e-10
    assert np.abs(got[:, :, 2] - direct[:, :, 2]).max() < tol_ll
    for c, o in enumerate(oracles):
        want = o.combined_greens(recalculate=recalculate)        # (M, 3, N, N)
        assert np.abs(got[:, c, :2] - want[:, :2]).max() < 1e-10, c
        assert np.abs(got[:, c, 2] - want[:, 2]).max() < tol_ll, c
    eff = [t for t in mcb.CombinedGreensIterator(mc, recalculate=recalculate, physical=False)]
    assert np.allclose(mc.hop[1] @ eff[3][1][0] @ mc.hop[0], got[3, 0, 1], atol=1e-12)
    mc.update(2)   # the sweep carries on
    assert mc.state()["n_updates"] == st0["n_updates"] + 2
    mc.close()


def test_combined_greens_iterator_state_checks():
    mc, oracles = _pair(4, 1.0, 1.0, 0.1, 5, 1)
    mc.build_stack()   # current_slice == M on the way down: not the measurement point
    with pytest.raises(RuntimeError, match="current_slice == 1"):
        next(iter(mcb.CombinedGreensIterator(mc)))
    mc.update(mc.M)
    it = iter(mcb.CombinedGreensIterator(mc))
    next(it)
    mc.update(1)       # the field may have changed: the iteration in progress is void
    with pytest.raises(RuntimeError, match="no iteration in progress"):
        next(it)
    mc.close()


def test_replay_of_recorded_configurations():
    """replay! (DQMC.jl:649-660) over ConfigRecorder data: bit-packed wire format round trip, and the batched
    measurement loop against one-at-a-time set_conf + calculate_greens(mc, 0) + measure."""
    model = mcb.HubbardModelAttractive(L=4, U=2.0)
    kw = dict(beta=1.0, delta_tau=0.1, safe_mult=5)
    mc = mcb.DQMC(model, n_chains=3, seed=9, **kw)
    N, M = mc.N, mc.M
    # wire format: compress == the library's packing == numpy's BitArray layout, decompress inverts it
    bits = mc.conf_bits()
    conf = mc.conf()
    for c in range(3):
        assert np.array_equal(bits[c], mcb.compress(conf[c]))
        assert np.array_equal(mcb.decompress(bits[c], N, M), conf[c])
    flipped = -conf
    mc.set_conf_bits(np.stack([mcb.compress(flipped[c]) for c in range(3)]))
    assert np.array_equal(mc.conf(), flipped)
    # 7 recorded configurations through a 3-chain handle: 2 full batches + a remainder of 1
    rng = np.random.default_rng(3)
    confs = rng.choice(np.array([-1, 1], dtype=np.int8), size=(7, N, M))
    packed = np.stack([mcb.compress(c) for c in confs])
    res = mc.replay(packed)
    assert [r["count"] for r in res] == [2, 1]
    one = mcb.DQMC(model, n_chains=1, seed=9, **kw)
    one.enable_measurements()
    per_conf = []
    for c in confs:
        one.L.mcb_dqmc_reset_observables(one.h)
        one.set_conf(c[None])
        G = one.greens_at(0)
        o = orc.OracleDQMC(N, M, 5, 0.1, 2.0, one.hop, c)
        assert relerr(G[0], o.greens_at(0)) < RTOL_G
        one.measure()
        per_conf.append(one.observables())
    for key in ("occupation", "double_occupancy", "e_kin", "e_int"):
        want_full = np.array([[per_conf[b * 3 + c][key][0] for b in range(2)] for c in range(3)]).mean(axis=1)
        assert np.allclose(res[0][key], want_full, rtol=1e-10, atol=1e-12), key
        assert np.allclose(res[1][key], per_conf[6][key], rtol=1e-10, atol=1e-12), key
    assert np.allclose(res[0]["cdc"][1], np.mean([per_conf[1]["cdc"][0], per_conf[4]["cdc"][0]], axis=0), atol=1e-10)
    mc.close(); one.close()


def test_checkpoint_and_resume_continue_the_same_chains():
    """save / resume! (FileIO.jl:38-156, DQMC.jl:517-521): 2 sweeps + checkpoint + resume in a NEW handle + 2 sweeps
    gives the chains of 4 uninterrupted sweeps (same random streams; G rebuilt from the stack at the resume)."""
    a, _ = _pair(4, 2.0, 1.0, 0.1, 5, 3, seed=21)
    b, _ = _pair(4, 2.0, 1.0, 0.1, 5, 3, seed=21)
    a.build_stack(); a.sweep(4)
    b.build_stack(); b.sweep(2)
    ck = b.checkpoint()
    assert ck["last_sweep"] == 2
    b.update(1)
    with pytest.raises(ValueError):
        b.checkpoint()
    b.close()
    c, _ = _pair(4, 2.0, 1.0, 0.1, 5, 3, seed=21)
    c.resume(ck)
    assert c.state()["n_updates"] == 2 * 2 * c.M
    c.sweep(2)
    assert a.state() == c.state()
    assert np.array_equal(a.conf(), c.conf())
    assert relerr(a.greens(False), c.greens(False)) < 1e-9
    a.close(); c.close()
