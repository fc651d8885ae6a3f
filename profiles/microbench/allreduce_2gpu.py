"""Two ranks, two GPUs: mcb_dqmc_allreduce_observables over NCCL (loaded by the library with dlopen) against the sum of
the two ranks' own moments. Launch: torchrun --nproc-per-node 2 profiles/microbench/allreduce_2gpu.py
(torch.distributed / gloo is used only to hand the NCCL unique id to rank 1 and to gather the reference values)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import montecarlo_jl_b200 as mcb  # noqa: E402
from montecarlo_jl_b200 import sharding  # noqa: E402

rank, local, world = sharding.dist_env()
dist.init_process_group("gloo")
uid = [sharding.NcclComm.unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
comm = sharding.NcclComm(uid[0], world, rank, local)
chains = 5 if rank == 0 else 3   # ragged shards
model = mcb.HubbardModelAttractive(L=4, U=2.0)
mc = mcb.DQMC(model, n_chains=chains, seed=5, chain_offset=0 if rank == 0 else 5, device=local, beta=1.0, delta_tau=0.1,
              safe_mult=5, thermalization=1, sweeps=4, measure_rate=1)
mc.run()
own = mc.allreduce_moments(None)
tot = mc.allreduce_moments(comm)
gathered = [None] * world
dist.all_gather_object(gathered, [np.asarray(x) for x in own])
for i in range(5):
    want = sum(np.asarray(g[i], dtype=np.float64) for g in gathered)
    assert np.allclose(np.asarray(tot[i], dtype=np.float64), want, rtol=1e-13, atol=1e-15), (rank, i)
# the time-displaced records travel the same way
mc.update(mc.M)
mc.measure_susceptibilities(recalculate=5)
own_td = mc.allreduce_susceptibility_moments(None)
tot_td = mc.allreduce_susceptibility_moments(comm)
gathered = [None] * world
dist.all_gather_object(gathered, [np.asarray(x) for x in own_td])
for i in range(3):
    want = sum(np.asarray(g[i], dtype=np.float64) for g in gathered)
    assert np.allclose(np.asarray(tot_td[i], dtype=np.float64), want, rtol=1e-13, atol=1e-15), (rank, "td", i)
if rank == 0:
    print(f"allreduce over {world} ranks ok: chains {tot[4]}, <n> = {tot[0][0] / tot[4]:.6f}; susceptibilities: chains {tot_td[2]}, cds[0] = {tot_td[0][0] / tot_td[2]:.6f}")
comm.close()
mc.close()
dist.destroy_process_group()
