"""GPU parity of DQMC{HubbardModelRepulsive} through the C ABI (mcb_dqmc_rep_*): same hopping
exponentials, HS field and Philox stream as the CPU oracle -> after every update() of a sweep both blocks of the
Green's function agree to 1e-10 relative max-norm and the fields are identical."""
import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
RTOL_G = 1e-10


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def _pair(L, U, beta, dtau, safe_mult, n_chains, seed=77, general=False, chain_offset=0):
    model = mcb.HubbardModelRepulsive(L=L, U=U)
    mc = mcb.DQMC(model, n_chains=n_chains, seed=seed, chain_offset=chain_offset, beta=beta, delta_tau=dtau,
                  safe_mult=safe_mult, general_path=general)
    conf0 = mc.conf()
    oracles = [orc.OracleDQMCRepulsive(L * L, mc.M, safe_mult, dtau, U, mc.hop, conf0[c], seed=seed,
                                       chain=chain_offset + c) for c in range(n_chains)]
    return mc, oracles


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("L,U,beta,safe_mult,chains", [(2, 1.0, 1.0, 5, 3), (4, 2.0, 1.0, 5, 3), (4, 4.0, 2.0, 10, 2),
                                                  (6, 4.0, 1.0, 5, 2)])
def test_every_update_of_a_sweep_matches_oracle(L, U, beta, safe_mult, chains, general):
    mc, oracles = _pair(L, U, beta, 0.1, safe_mult, chains, general=general)
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    G = mc.greens(physical=False)
    assert G.shape == (chains, 2, L * L, L * L)
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
            assert relerr(G[c, 0], o.greens()[0]) < RTOL_G, (u, c, "up")
            assert relerr(G[c, 1], o.greens()[1]) < RTOL_G, (u, c, "down")
            assert np.array_equal(conf[c], o.conf()), (u, c)
    stats = mc.stats()
    for c, o in enumerate(oracles):
        so = o.state()
        assert stats["acc_local"][c] == so["acc_local"] and stats["prop_local"][c] == so["prop_local"]
        assert stats["n_nonfinite"][c] == 0
    Gp = mc.greens(physical=True)
    for c, o in enumerate(oracles):
        assert relerr(Gp[c], o.greens(physical=True)) < RTOL_G
    mc.close()


def test_8x8_beta4_partial_sweep_and_from_scratch_greens():
    """N = 64 at U = 4: a stretch of updates across a stabilisation point, then calculate_greens(mc, slice) of both
    blocks against the oracle."""
    mc, oracles = _pair(8, 4.0, 4.0, 0.1, 10, 2)
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    mc.update(23)
    G = mc.greens(physical=False)
    conf = mc.conf()
    for c, o in enumerate(oracles):
        o.update(23)
        assert relerr(G[c], o.greens()) < RTOL_G
        assert np.array_equal(conf[c], o.conf())
    for s in (0, 7, 40):
        Gs = mc.greens_at(s)
        for c, o in enumerate(oracles):
            assert relerr(Gs[c], o.greens_at(s)) < 1e-9
    mc.close()


def test_batched_updates_equal_single_updates_and_chain_ids_are_global():
    a, _ = _pair(4, 4.0, 1.0, 0.1, 5, 4)
    b, _ = _pair(4, 4.0, 1.0, 0.1, 5, 2, chain_offset=2)
    a.build_stack(); b.build_stack()
    a.update(13)
    for _ in range(13):
        b.update(1)
    assert np.array_equal(a.conf()[2:], b.conf())
    assert np.abs(a.greens(False)[2:] - b.greens(False)).max() < 1e-12
    a.close(); b.close()


def test_replay_and_run_schedule_of_the_two_block_handle():
    """replay! (DQMC.jl:649-660) and the measurement schedule of run! for HubbardModelRepulsive: the batched replay of
    recorded fields equals measuring the from-scratch G of each field, which the oracle's general kernels reproduce."""
    model = mcb.HubbardModelRepulsive(L=4, U=2.0)
    kw = dict(beta=1.0, delta_tau=0.1, safe_mult=5)
    mc = mcb.DQMC(model, n_chains=2, seed=4, **kw)
    rng = np.random.default_rng(8)
    confs = rng.choice(np.array([-1, 1], dtype=np.int8), size=(4, 16, 10))
    res = mc.replay(np.stack([mcb.compress(c) for c in confs]))
    assert [r["count"] for r in res] == [2]
    tables = orc.MeasurementTables(4, 5)
    for c in range(2):
        want = []
        for conf in (confs[c], confs[2 + c]):
            o = orc.OracleDQMCRepulsive(16, 10, 5, 0.1, 2.0, mc.hop, conf)
            g = [mc.hop[1] @ gb @ mc.hop[0] for gb in o.greens_at(0)]
            G2 = np.zeros((32, 32))
            G2[:16, :16], G2[16:, 16:] = g
            want.append(orc.measure(G2, tables, flavors=2))
        for k in ("cdc", "sdc_x", "sdc_z", "pc", "mz", "occupation"):
            assert np.abs(res[0][k][c] - np.mean([w[k] for w in want], axis=0)).max() < 1e-9, (k, c)
    mc.close()
    # run!: measurements every 2nd sweep after 2 thermalisation sweeps
    mc = mcb.DQMC(model, n_chains=2, seed=4, thermalization=2, sweeps=6, measure_rate=2, **kw)
    conf0 = mc.conf()
    obs = mc.run()
    for c in range(2):
        o = orc.OracleDQMCRepulsive(16, 10, 5, 0.1, 2.0, mc.hop, conf0[c], seed=4, chain=c)
        o.build_stack()
        want = orc.run_measured(o, tables, thermalization=2, sweeps=6, measure_rate=2)
        assert want["count"] == obs["count"] == 3
        G2 = np.zeros((32, 32))
        G2[:16, :16], G2[16:, 16:] = obs["G"][c]
        assert np.abs(G2 - want["G"]).max() < 1e-9
        for k in ("cdc", "sdc_x", "sdc_z", "pc", "mz"):
            assert np.abs(obs[k][c] - want[k]).max() < 1e-9, (k, c)
    mc.close()


def test_unequal_time_greens_of_both_blocks():
    """greens(mc, k, l) per block against the oracle's UnequalTimeStack restatement; sweep state untouched."""
    mc, oracles = _pair(4, 4.0, 1.0, 0.1, 5, 2)
    mc.build_stack()
    for o in oracles:
        o.build_stack()
    mc.update(3)
    for o in oracles:
        o.update(3)
    G0 = mc.greens(physical=False)
    for k, l in ((0, 0), (7, 2), (2, 7), (10, 4), (5, 10)):
        G = mc.greens_unequal(k, l, physical=False)
        assert G.shape == (2, 2, 16, 16)
        for c, o in enumerate(oracles):
            assert relerr(G[c, 0], o.up.greens_unequal(k, l)) < 1e-9, (k, l, c)
            assert relerr(G[c, 1], o.dn.greens_unequal(k, l)) < 1e-9, (k, l, c)
    assert np.array_equal(mc.greens(physical=False), G0)
    mc.close()


def test_combined_greens_iterator_of_both_blocks():
    """CombinedGreensIterator (unequal_time_stack.jl:746-885) per block against greens(mc, k, l) and the CPU restatement."""
    mc, oracles = _pair(4, 4.0, 2.0, 0.1, 5, 2)
    mc.build_stack()
    mc.update(mc.M)
    for o in oracles:
        o.build_stack()
        o.update(mc.M)
    assert mc.state()["current_slice"] == 1 and mc.state()["direction"] == 1
    # recalculate = 10: at U = 4 the default (20 = M, i.e. never) lets the reference's own scheme drift to 3e-7 in G(l <- l)
    got = [t for t in mcb.CombinedGreensIterator(mc, recalculate=10)]
    assert len(got) == mc.M and got[0][0].shape == (2, 2, 16, 16)
    for l in (1, 4, 5, 6, 9, 10, 11, 19, 20):
        G0l, Gl0, Gll = got[l - 1]
        assert np.abs(G0l - mc.greens_unequal(0, l)).max() < 1e-10, l
        assert np.abs(Gl0 - mc.greens_unequal(l, 0)).max() < 1e-10, l
        assert np.abs(Gll - mc.greens_unequal(l, l)).max() < 1e-8, l
    for c, o in enumerate(oracles):
        for f, blk in enumerate((o.up, o.dn)):
            want = blk.combined_greens(recalculate=10)
            have = np.stack([np.stack([t[w][c, f] for w in range(3)]) for t in got])
            assert np.abs(have[:, :2] - want[:, :2]).max() < 1e-10, (c, f)
            assert np.abs(have[:, 2] - want[:, 2]).max() < 1e-8, (c, f)
    mc.close()


@pytest.mark.parametrize("L,U,beta,K", [(2, 1.0, 1.0, 4), (4, 4.0, 2.0, 5)])
def test_susceptibilities_of_the_two_block_handle(L, U, beta, K):
    """Charge / spin density and pairing susceptibility (measurements.jl:56-58, 144-146, 207) with the reference's general
    two-flavor kernels for packed Green's functions, accumulated on the device from the iterators of both blocks, against
    the oracle's literal evaluation on the 2N x 2N matrices (pinned against exact diagonalisation on the CPU,
    tests/test_oracle_measure.py). sds_x couples the two blocks."""
    mc, oracles = _pair(L, U, beta, 0.1, 5, 2)
    mc.enable_measurements(K)
    mc.build_stack()
    mc.update(mc.M)
    for o in oracles:
        o.build_stack()
        o.update(mc.M)
    tables = orc.MeasurementTables(L, K)
    G0, st0 = mc.greens(physical=False), mc.state()
    mc.measure_susceptibilities(recalculate=10)
    assert np.array_equal(G0, mc.greens(physical=False)) and st0 == mc.state()
    got = mc.susceptibilities()
    assert got["count"] == 1
    for c, o in enumerate(oracles):
        want = orc.susceptibilities(o, tables, recalculate=10, T=mc.model.hopping_matrix())
        for k in ("cds", "sds_x", "sds_y", "sds_z", "ps", "ccs"):
            scale = max(1.0, np.abs(want[k]).max())
            assert np.abs(got[k][c] - want[k]).max() < 1e-8 * scale, (c, k)
        assert np.abs(want["sds_x"]).max() > 0.05 and np.abs(want["sds_x"] - want["sds_z"]).max() > 1e-6
    mc.close()


def test_scheduled_susceptibilities_of_the_two_block_handle():
    """With schedule_susceptibilities the time-displaced group is measured wherever the equal-time one is (run!,
    DQMC.jl:423-438): same sums as the same measurements taken by hand."""
    def make():
        mc, _ = _pair(4, 4.0, 1.0, 0.1, 5, 2)
        mc.enable_measurements(5)
        mc.build_stack()
        mc.update(mc.M)
        return mc
    a = make()
    a.schedule_susceptibilities(True, recalculate=5)
    a.sweep(4, measure_every=2)
    b = make()
    for k in range(1, 5):
        b.sweep(1)
        if (k + 1) % 2 == 0:          # the handle started half a sweep in (see tests/test_gpu_measure.py)
            b.measure()
            b.measure_susceptibilities(recalculate=5)
    sa, sb = a.susceptibilities(), b.susceptibilities()
    assert sa["count"] == sb["count"] == 2
    for k in ("cds", "sds_x", "sds_z", "ps", "ccs"):
        assert np.array_equal(sa[k], sb[k]), k
    a.close()
    b.close()


def test_replay_takes_the_time_displaced_group_along():
    """replay! (DQMC.jl:649-660) of the two-block handle with the time-displaced group scheduled: per recorded
    configuration the iterators start from G(0) of both blocks. Reference: the general kernels on the stabilised
    greens(mc, k, l) of every slice."""
    L, U, K = 2, 1.0, 4
    mc, _ = _pair(L, U, 1.0, 0.1, 5, 2)
    mc.enable_measurements(K)
    mc.schedule_susceptibilities(True, recalculate=5)
    N, M = mc.N, mc.M
    confs = np.random.default_rng(8).choice(np.array([-1, 1], dtype=np.int8), size=(4, N, M))
    res = mc.replay(np.stack([mcb.compress(c) for c in confs]))
    assert res[0]["count"] == 2
    got = mc.susceptibilities()
    assert got["count"] == 2
    tables = orc.MeasurementTables(L, K)
    one, _ = _pair(L, U, 1.0, 0.1, 5, 1)
    want = []
    for c in confs:
        one.set_conf(c[None])
        one.greens_at(0)
        G00 = one.greens(physical=True)[0]                       # (2, N, N)
        tri = [[one.greens_unequal(0, l)[0], one.greens_unequal(l, 0)[0], one.greens_unequal(l, l)[0]] for l in range(1, M + 1)]
        cu = np.stack([np.stack([t[w][0] for w in range(3)]) for t in tri])
        cd = np.stack([np.stack([t[w][1] for w in range(3)]) for t in tri])
        want.append(orc.susceptibilities(dict(G00=G00, cu=cu, cd=cd, dtau=0.1), tables, T=mc.model.hopping_matrix()))
    for slot in range(2):
        for k in ("cds", "sds_x", "sds_z", "ps", "ccs"):
            w = 0.5 * (want[slot][k] + want[2 + slot][k])
            assert np.abs(got[k][slot] - w).max() < 1e-7 * max(1.0, np.abs(w).max()), (slot, k)
    mc.close(); one.close()


def test_checkpoint_and_resume_continue_the_same_chains():
    """save / resume! (FileIO.jl:38-156, DQMC.jl:517-521) for the two-block handle: a run that is stopped after two
    sweeps, saved as bit-packed field + sweep count and resumed in a new handle ends where the uninterrupted run does."""
    full, _ = _pair(4, 4.0, 1.0, 0.1, 5, 3)
    full.build_stack()
    full.sweep(4)
    first, _ = _pair(4, 4.0, 1.0, 0.1, 5, 3)
    first.build_stack()
    first.sweep(2)
    ck = first.checkpoint()
    assert np.array_equal(mcb.decompress(ck["conf_bits"][1], 16, first.M), first.conf()[1])
    first.close()
    second, _ = _pair(4, 4.0, 1.0, 0.1, 5, 3)
    second.resume(ck)
    second.sweep(2)
    assert np.array_equal(second.conf(), full.conf())
    # the resumed stack is rebuilt from scratch: agreement to stabilisation accuracy, not bitwise
    assert np.abs(second.greens(False) - full.greens(False)).max() < 1e-8
    full.close(); second.close()
