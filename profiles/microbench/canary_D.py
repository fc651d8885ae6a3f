"""Propagation-error canary (stack.jl:540-550) of config D over one sweep: per-chain maxima. Used to compare builds / switches."""
import sys
import numpy as np
sys.path.insert(0, ".")
import montecarlo_jl_b200 as mcb

sweeps = int(sys.argv[1]) if len(sys.argv) > 1 else 1
mc = mcb.DQMC(mcb.HubbardModelAttractive(L=20, U=4.0), n_chains=32, seed=123, beta=16.0, delta_tau=0.05, safe_mult=10,
              check_propagation_error=True)
mc.build_stack()
mc.update(21)
for s in range(sweeps):
    mc.sweep(1)
    err = mc.stats()["max_propagation_error"]
    print(f"sweep {s + 1}: max {err.max():.3e} (chain {err.argmax()}), median {np.median(err):.3e}, sorted top 4 {np.sort(err)[-4:]}", flush=True)
G = mc.greens(physical=False)
print("checksum", float(np.abs(G).sum()), "conf sum", int(mc.conf().astype(np.int64).sum()))
mc.close()
