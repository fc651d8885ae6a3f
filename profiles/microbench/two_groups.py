"""Experiment: does a batch of chains run faster as G independent handles driven from G host threads (each with its own
stream, so that the latency-bound factorisations of one group overlap the pipe-bound products of another) than as one
handle? Prints wall-clock chain-sweeps/s for G = 1, 2, 4 at a BASELINE workload.

    python profiles/microbench/two_groups.py dqmc_D32 3
"""
import sys
import threading
import time

sys.path.insert(0, ".")
import torch  # noqa: E402  (device bring-up)

import montecarlo_jl_b200 as mcb  # noqa: E402
from bench import WORKLOADS, SEED  # noqa: E402


def run(name, groups, sweeps, warmup=2):
    w = WORKLOADS[name]
    per = w["chains"] // groups
    model = mcb.HubbardModelAttractive(L=w["L"], U=w["U"])
    mcs = [mcb.DQMC(model, n_chains=per, seed=SEED, chain_offset=g * per, device=0, beta=w["beta"], delta_tau=w["dtau"],
                    safe_mult=w["safe_mult"], check_propagation_error=False) for g in range(groups)]
    for mc in mcs:
        mc.build_stack()

    def work(mc, n, stagger):
        # a stagger of half a stabilisation period puts the groups into different kernel classes at the same time
        for _ in range(stagger):
            mc.update(1)
        mc.sweep(n)

    def timed(n, stagger):
        ths = [threading.Thread(target=work, args=(mc, n, (stagger * g) if stagger else 0)) for g, mc in enumerate(mcs)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    timed(warmup, w["safe_mult"] // max(1, groups))
    dt = timed(sweeps, 0)
    val = w["chains"] * sweeps / dt
    for mc in mcs:
        mc.close()
    return val, dt / sweeps * 1e3


if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "dqmc_D32"
    sweeps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    for g in (1, 2, 4):
        v, ms = run(name, g, sweeps)
        print(f"{name} groups {g}: {v:.2f} chain-sweeps/s ({ms:.1f} ms per sweep, wall clock)", flush=True)
