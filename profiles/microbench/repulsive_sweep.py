"""Sweep rate of DQMC{HubbardModelRepulsive} through the C ABI (mcb_dqmc_rep_*): 8x8, U = 4, beta = 8,
dtau = 0.1, safe_mult = 10, 512 chains on one GPU; device time of the library's own CUDA events."""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import montecarlo_jl_b200 as mcb

chains = int(sys.argv[1]) if len(sys.argv) > 1 else 512
sweeps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
model = mcb.HubbardModelRepulsive(L=8, U=4.0)
mc = mcb.DQMC(model, n_chains=chains, seed=123, beta=8.0, delta_tau=0.1, safe_mult=10)
mc.build_stack()
mc.sweep(1)                                   # warm-up
mc.sweep(sweeps)
ms = mc.L.mcb_dqmc_rep_last_elapsed_ms(mc.h)
st = mc.stats()
print(json.dumps({"workload": "dqmc_repulsive_8x8_U4_beta8", "chains": chains, "sweeps": sweeps,
                  "ms_per_sweep": ms / sweeps, "chain_sweeps_per_s": chains * sweeps / (ms * 1e-3),
                  "acceptance": float(st["acc_local"].sum() / st["prop_local"].sum()),
                  "nonfinite": int(st["n_nonfinite"].sum())}))
mc.close()
