"""The observable all-reduce behind the C ABI (mcb_comm_*, mcb_dqmc_allreduce_observables).
On the one-GPU test box: the moments of one rank against numpy, without a communicator and through a one-rank NCCL
communicator (dlopen of libnccl, ncclCommInitRank, the device-side moment kernel, the staging buffers). The two-rank
run is profiles/microbench/allreduce_2gpu.py (gpurun --gpus 2), recorded under profiles/."""
import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import sharding

pytestmark = pytest.mark.gpu


def _measured_handle(chains=6, offset=0):
    model = mcb.HubbardModelAttractive(L=4, U=2.0)
    mc = mcb.DQMC(model, n_chains=chains, seed=5, chain_offset=offset, beta=1.0, delta_tau=0.1, safe_mult=5,
                  thermalization=1, sweeps=4, measure_rate=1)
    mc.run()
    return mc


def _numpy_moments(mc):
    n = mc._f("observable_len")(mc.h)
    import ctypes as C
    buf = np.empty(mc.n_chains * n)
    cnt = C.c_int64()
    mcb._lib.check(mc._f("get_observables")(mc.h, buf, C.byref(cnt)))
    rec = buf.reshape(mc.n_chains, n) / cnt.value
    G = mc.observables()["G"]
    return rec.sum(0), (rec * rec).sum(0), G.sum(0), (G * G).sum(0)


@pytest.mark.parametrize("with_comm", [False, True])
def test_moments_of_one_rank(with_comm):
    mc = _measured_handle()
    comm = sharding.NcclComm(sharding.NcclComm.unique_id(), 1, 0, 0) if with_comm else None
    s, s2, g, g2, chains = mc.allreduce_moments(comm)
    ws, ws2, wg, wg2 = _numpy_moments(mc)
    assert chains == 6
    assert np.allclose(s, ws, rtol=1e-13, atol=1e-15) and np.allclose(s2, ws2, rtol=1e-13, atol=1e-15)
    assert np.allclose(g, wg, rtol=1e-13, atol=1e-15) and np.allclose(g2, wg2, rtol=1e-13, atol=1e-15)
    # the same numbers as the torch.distributed path of the Python host
    obs = mc.observables()
    mom, cnt = sharding.local_moments({k: v for k, v in obs.items() if k in ("n", "double_occupancy")})
    assert np.isclose(mom["n"][0], s[0]) and np.isclose(mom["double_occupancy"][1], s2[1]) and cnt == chains
    if comm is not None:
        comm.close()
    mc.close()


@pytest.mark.parametrize("with_comm", [False, True])
def test_susceptibility_moments_of_one_rank(with_comm):
    import ctypes as C
    mc = _measured_handle(chains=4)
    s, s2, chains = mc.allreduce_susceptibility_moments()
    assert chains == 0 and not s.any()                      # nothing time-displaced measured yet
    mc.update(mc.M)                                         # to the measurement point
    for _ in range(2):
        mc.measure_susceptibilities(recalculate=5)
        mc.sweep(1)
    comm = sharding.NcclComm(sharding.NcclComm.unique_id(), 1, 0, 0) if with_comm else None
    s, s2, chains = mc.allreduce_susceptibility_moments(comm)
    n = mc.L.mcb_dqmc_susceptibility_len(mc.h)
    buf, cnt = np.empty(mc.n_chains * n), C.c_int64()
    mcb._lib.check(mc.L.mcb_dqmc_get_susceptibilities(mc.h, buf, C.byref(cnt)))
    rec = buf.reshape(mc.n_chains, n) / cnt.value
    assert chains == 4 and cnt.value == 2
    assert np.allclose(s, rec.sum(0), rtol=1e-13, atol=1e-15) and np.allclose(s2, (rec * rec).sum(0), rtol=1e-13, atol=1e-15)
    if comm is not None:
        comm.close()
    mc.close()


def test_rank_without_measurements_contributes_nothing():
    model = mcb.HubbardModelAttractive(L=4, U=2.0)
    mc = mcb.DQMC(model, n_chains=2, seed=5, beta=1.0, delta_tau=0.1, safe_mult=5)
    mc.enable_measurements()
    s, s2, g, g2, chains = mc.allreduce_moments()
    assert chains == 0 and not s.any() and not g2.any()
    mc.close()


def test_errors():
    model = mcb.HubbardModelAttractive(L=4, U=2.0)
    mc = mcb.DQMC(model, n_chains=2, seed=5, beta=1.0, delta_tau=0.1, safe_mult=5)
    with pytest.raises(RuntimeError):
        mc.allreduce_moments()          # measurement tables not set
    with pytest.raises(Exception):
        sharding.NcclComm(b"\0" * 128, 2, 5, 0)   # rank out of range
    mc.close()
