"""State-level parity on the code paths that bench.py times for BASELINE.json configs[2] and configs[3]: the SAME
chain counts per GPU (128 at 12 x 12, 32 at 20 x 20) and the real beta / dtau, because kernel selection depends on
the batch size (whole-matrix GEMM, CTAs per chain of the delayed sweep, cluster UDT waves).
Tolerance: max|G_gpu - G_oracle| / max|G_oracle| <= 1e-10 (north_star) with identical HS fields.
"""
import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import _lib
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
RTOL_G = 1e-10


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def _cm(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).transpose(0, 2, 1)).reshape(-1)


def _uncm(buf, b, n):
    return buf.reshape(b, n, n).transpose(0, 2, 1).copy()


def _run_against_oracle(L, U, beta, dtau, n_chains, watch, checkpoints, seed=123, check=False):
    model = mcb.HubbardModelAttractive(L=L, U=U)
    mc = mcb.DQMC(model, n_chains=n_chains, seed=seed, beta=beta, delta_tau=dtau, safe_mult=10,
                  check_propagation_error=check)
    conf0 = mc.conf()
    oracles = {c: orc.OracleDQMC(L * L, mc.M, 10, dtau, U, mc.hop, conf0[c], seed=seed, chain=c,
                                 check_propagation=check) for c in watch}
    mc.build_stack()
    for o in oracles.values():
        o.build_stack()
    G = mc.greens(physical=False)
    for c, o in oracles.items():
        assert relerr(G[c], o.greens()) < RTOL_G, ("build_stack", c)
    done = 0
    for upto in checkpoints:
        mc.update(upto - done)
        G = mc.greens(physical=False)
        conf = mc.conf()
        st = mc.state()
        for c, o in oracles.items():
            o.update(upto - done)
            so = o.state()
            assert (st["current_slice"], st["direction"]) == (so["current_slice"], so["direction"])
            assert np.array_equal(conf[c], o.conf()), (upto, c)
            assert relerr(G[c], o.greens()) < RTOL_G, (upto, c)
        done = upto
    stats = mc.stats()
    for c, o in oracles.items():
        assert stats["acc_local"][c] == o.state()["acc_local"]
        assert stats["n_nonfinite"][c] == 0
    return mc, oracles


def test_config_C_128_chains_as_benched():
    """12 x 12, U = 4, beta = 10, dtau = 0.1, 128 chains on the GPU: chains 0 / 63 / 127 after updates 1, 10, 11
    (first stabilisation), 25, 100, 101 (turn-around at slice 1) and 115."""
    mc, _ = _run_against_oracle(12, 4.0, 10.0, 0.1, 128, watch=(0, 63, 127), checkpoints=(1, 10, 11, 25, 100, 101, 115))
    mc.close()


def test_config_C_propagation_canary_on_the_resident_path():
    """The same lattice with check_propagation_error on (stack.jl:540-550): the wrapped G of the chain-resident kernel
    against the freshly stabilised one at every stabilisation of a full sweep, and the oracle's own canary for one chain."""
    mc, oracles = _run_against_oracle(12, 4.0, 10.0, 0.1, 128, watch=(5,), checkpoints=(10, 11, 21), check=True)
    mc.sweep(1)
    for o in oracles.values():
        o.sweep(1)
    err = mc.stats()["max_propagation_error"]
    assert np.all(np.isfinite(err)) and err.max() < 1e-6, err.max()
    want = oracles[5].state()["max_prop_err"]
    # (rounding noise amplified by ten wraps: the two implementations agree in magnitude, not in digits)
    assert want < 1e-6 and err[5] < 10 * want + 1e-8, (err[5], want)
    assert np.all(mc.stats()["n_nonfinite"] == 0)
    mc.close()


def test_config_D_32_chains_real_beta():
    """20 x 20, U = 4, beta = 16, dtau = 0.05 (M = 320, 33 stack slots, singular values spanning tens of decades),
    32 chains, propagation check on: build_stack + the first 21 updates (two stabilisations) of chains 0 / 31, then
    the propagation canary of the reference (stack.jl:540-550) over one full GPU sweep."""
    mc, oracles = _run_against_oracle(20, 4.0, 16.0, 0.05, 32, watch=(0, 31), checkpoints=(1, 10, 11, 21), check=True)
    mc.sweep(1)
    err = mc.stats()["max_propagation_error"]
    assert np.all(np.isfinite(err)) and err.max() < 1e-6, err.max()
    assert np.all(mc.stats()["n_nonfinite"] == 0)
    mc.close()


@pytest.mark.parametrize("n,batch", [(144, 96), (144, 3), (400, 2)])
def test_calculate_greens_from_udts_large(n, batch):
    """calculate_greens_AVX! (stack.jl:337-393) at the sizes of configs[2] / configs[3]; batch = 96 takes the
    whole-matrix GEMM with fused row / column / diagonal operands."""
    rng = np.random.default_rng(n + batch)
    uniq = min(batch, 3)
    trip = []
    for side in range(2):
        Us, Ds, Ts = [], [], []
        for k in range(uniq):
            X = rng.standard_normal((n, n)) @ np.diag(np.logspace(2, -8, n))
            U, D, T, _ = orc.udt_pivot(X, True)
            Us.append(U); Ds.append(D); Ts.append(T)
        idx = np.arange(batch) % uniq
        trip.append((np.array(Us)[idx], np.array(Ds)[idx], np.array(Ts)[idx]))
    (Ul, Dl, Tl), (Ur, Dr, Tr) = trip
    L = _lib.lib()
    G = np.empty(batch * n * n)
    _lib.check(L.mcb_linalg_calculate_greens(n, batch, _cm(Ul), np.ascontiguousarray(Dl).reshape(-1), _cm(Tl),
                                             _cm(Ur), np.ascontiguousarray(Dr).reshape(-1), _cm(Tr), G, 0))
    G = _uncm(G, batch, n)
    for k in range(uniq):
        want = orc.calculate_greens_udt(Ul[k], Dl[k], Tl[k], Ur[k], Dr[k], Tr[k])
        for b in range(k, batch, uniq):
            assert relerr(G[b], want) < RTOL_G, (k, b)


@pytest.mark.parametrize("L,beta,dtau,chains", [(12, 2.0, 0.1, 2), (20, 1.0, 0.05, 2)])
def test_unequal_time_greens_large(L, beta, dtau, chains):
    """greens(mc, k, l) (unequal_time_stack.jl:275-305) at N = 144 / 400 against the oracle."""
    model = mcb.HubbardModelAttractive(L=L, U=4.0)
    mc = mcb.DQMC(model, n_chains=chains, seed=5, beta=beta, delta_tau=dtau, safe_mult=10)
    conf0 = mc.conf()
    o = orc.OracleDQMC(L * L, mc.M, 10, dtau, 4.0, mc.hop, conf0[1], seed=5, chain=1)
    M = mc.M
    for (k, l) in ((M, 0), (0, M // 2), (M // 2 + 3, 3), (7, 7)):
        G = mc.greens_unequal(k, l, physical=False)
        assert relerr(G[1], o.greens_unequal(k, l)) < 1e-9, (k, l)
    mc.close()


def test_two_handles_from_two_host_threads():
    """Handles are independent: two of them driven from two host threads (each on its own stream, both going through the
    two-stream blocked UDT, whose auxiliary streams and events are per caller stream) give what each gives alone."""
    import threading
    L, beta, dtau, chains, n_upd = 16, 1.0, 0.1, 8, 12

    def make(offset):
        mc = mcb.DQMC(mcb.HubbardModelAttractive(L=L, U=4.0), n_chains=chains, seed=123, chain_offset=offset, beta=beta,
                      delta_tau=dtau, safe_mult=5, check_propagation_error=False)
        mc.build_stack()
        return mc

    alone = []
    for off in (0, chains):
        mc = make(off)
        mc.update(n_upd)
        alone.append((mc.greens(physical=False), mc.conf()))
        mc.close()
    mcs = [make(0), make(chains)]
    errs = []

    def work(mc):
        try:
            mc.update(n_upd)
        except Exception as e:   # surfaced below: an exception in a thread would otherwise pass silently
            errs.append(e)

    ths = [threading.Thread(target=work, args=(mc,)) for mc in mcs]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
    for mc, (G, conf) in zip(mcs, alone):
        assert np.array_equal(mc.conf(), conf)
        assert relerr(mc.greens(physical=False), G) <= 1e-12
        mc.close()
