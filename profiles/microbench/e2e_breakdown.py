"""Wall-clock breakdown of bench.py's e2e step (set_conf, build_stack, sweep, get_greens, get_conf) per call."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import _lib
from bench import WORKLOADS, SEED

name = sys.argv[1] if len(sys.argv) > 1 else "dqmc_C128"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
w = WORKLOADS[name]
mc = mcb.DQMC(mcb.HubbardModelAttractive(L=w["L"], U=w["U"]), n_chains=w["chains"], seed=SEED, device=0, beta=w["beta"],
              delta_tau=w["dtau"], safe_mult=w["safe_mult"], check_propagation_error=False)
L = mc.L
mc.build_stack()
mc.sweep(3)
conf_host = torch.empty(w["chains"] * mc.M * mc.N, dtype=torch.int8).pin_memory()
g_host = torch.empty(w["chains"] * mc.N * mc.N, dtype=torch.float64).pin_memory()
conf_np, g_np = conf_host.numpy(), g_host.numpy()
_lib.check(L.mcb_dqmc_get_conf(mc.h, conf_np))
calls = [("set_conf", lambda: L.mcb_dqmc_set_conf(mc.h, conf_np)), ("build_stack", lambda: L.mcb_dqmc_build_stack(mc.h)),
         ("sweep", lambda: L.mcb_dqmc_sweep(mc.h, 1, 0)), ("get_greens", lambda: L.mcb_dqmc_get_greens(mc.h, g_np, 0)),
         ("get_conf", lambda: L.mcb_dqmc_get_conf(mc.h, conf_np))]
for s in range(steps):
    out = []
    for nm, f in calls:
        t0 = time.perf_counter()
        _lib.check(f())
        out.append(f"{nm} {1e3 * (time.perf_counter() - t0):7.2f}")
    print(f"step {s}: " + " | ".join(out) + f" | device-timed sweep {L.mcb_dqmc_last_elapsed_ms(mc.h):.2f}", flush=True)
mc.close()
