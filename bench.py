#!/usr/bin/env python
"""bench.py — DQMC chain-sweeps/sec on B200 (BASELINE.json metric), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one full DQMC sweep (2*M update() calls = propagate + sweep_spatial per time slice,
DQMC.jl:420-438) of every chain in the batch. Workloads (BASELINE.json configs):
    dqmc_A512  8x8 attractive Hubbard, U=1, beta=8, dtau=0.1, safe_mult=10, 512 chains per GPU (default,
               configs[1], the configuration the metric is quoted on for one GPU)
    dqmc_C128  12x12, U=4, beta=10, 128 chains per GPU        (configs[2] sharded 1024/8)
    dqmc_D32   20x20, U=4, beta=16, dtau=0.05, 32 chains per GPU (configs[3] sharded 256/8)
    ising64    Ising L=4096 at T_c, 8 replicas per GPU          (configs[4] sharded 64/8)
Chains are independent, so N GPUs run N shards with no data-path collective ("scaling": "weak");
Philox streams are keyed by the global chain id.

Prints ONE JSON line (rank 0). `value` is device-resident throughput (CUDA events on the library's
stream, max over ranks); `e2e` runs the same sweep through the C ABI from pinned HOST buffers
(upload the HS field, rebuild the stack, sweep, download G and the field inside the timed region).
`--impl reference` times the CPU restatement of the reference (oracle/, one chain per host core).
"""
import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "dqmc_A512": dict(kind="dqmc", L=8, U=1.0, beta=8.0, dtau=0.1, safe_mult=10, chains=512),
    "dqmc_C128": dict(kind="dqmc", L=12, U=4.0, beta=10.0, dtau=0.1, safe_mult=10, chains=128),
    "dqmc_D32": dict(kind="dqmc", L=20, U=4.0, beta=16.0, dtau=0.05, safe_mult=10, chains=32),
    "ising64": dict(kind="ising", L=4096, beta=0.5 * math.log(1.0 + math.sqrt(2.0)), replicas=8),
}
SEED = 123


# --------------------------------------------------------------------------- helpers
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.device}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for k, name in enumerate(names):
                if len(r) > 4 + k and r[4 + k].lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "MEASURED_PEAKS.json"
    return 6650.0, "fallback (B200_PROFILING.md)"


def fp64_peak_tflops(torch, dev):
    """FP64 roofline denominator: MEASURED_PEAKS.json has no FP64 entry, so a cuBLAS DGEMM
    (4096^3, best of 5, CUDA events) is measured in this run, the same way the bf16 entry was."""
    n = 4096
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    torch.matmul(a, b)
    torch.cuda.synchronize(dev)
    best = float("inf")
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize(dev)
        best = min(best, e0.elapsed_time(e1))
    del a, b
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def dist_env():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def traffic_for(workload, kernel):
    """DRAM bytes per launch from an ncu --set full capture of THIS workload's kernel shape, else None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        return json.load(open(p)).get(workload, {}).get(kernel)
    return None


def config_for(name, w, world):
    """The `config` object of the JSON line. A function of the workload and the GPU count only, so that the
    reference arm and ours print the same keys and values (the driver compares them)."""
    if w["kind"] == "ising":
        return {"workload": name, "L": w["L"], "beta": w["beta"], "replicas_per_gpu": w["replicas"],
                "total_replicas": world * w["replicas"], "sweeps_per_step": 10,
                "l2_policy": "planes (2 MiB/replica) are L2-resident by design; HBM figure is algorithmic"}
    N, M, C_ = w["L"] ** 2, int(round(w["beta"] / w["dtau"])), w["chains"]
    big = C_ * N * N * 8 * (2 * (M // w["safe_mult"] + 2) + 11) > 126e6
    return {"workload": name, "lattice": f"{w['L']}x{w['L']}", "U": w["U"], "beta": w["beta"],
            "delta_tau": w["dtau"], "slices": M, "safe_mult": w["safe_mult"], "chains_per_gpu": C_,
            "total_chains": world * C_,
            "l2_policy": "working set (stack + work matrices) exceeds L2" if big else
                         "working set fits L2: state is device-resident by design, no flush"}


# --------------------------------------------------------------------------- CPU baseline (oracle)
def cpu_dqmc_baseline(w, target_seconds=15.0, threads=None):
    """Oracle (C++ restatement of the reference's loop nests) timed on the host: one chain per
    core, whole sweeps. Returns dict(value chain-sweeps/s, cores, sample)."""
    from oracle import oracle as orc
    import montecarlo_jl_b200 as mcb
    threads = threads or os.cpu_count() or 1
    model = mcb.HubbardModelAttractive(L=w["L"], U=w["U"])
    hop = model.hopping_exponentials(w["dtau"])
    M = int(round(w["beta"] / w["dtau"]))
    N = w["L"] ** 2
    try:
        orc.lib(native=True)
        native = True
    except Exception:
        native = False
    chains = [orc.OracleDQMC(N, M, w["safe_mult"], w["dtau"], w["U"], hop, model.rand_conf(M, SEED, c), seed=SEED,
                             chain=c, native=native) for c in range(threads)]
    for c in chains:
        c.build_stack()
    t0 = time.perf_counter()
    orc.sweep_many(chains[:1], 1, 1)  # calibrate (also the warm-up)
    one = time.perf_counter() - t0
    sweeps = max(1, int(target_seconds / max(one, 1e-3)))
    sweeps = min(sweeps, 50)
    t0 = time.perf_counter()
    orc.sweep_many(chains, sweeps, threads)
    dt = time.perf_counter() - t0
    return dict(value=threads * sweeps / dt, unit="chain-sweeps/s", cores=threads, kind="port",
                sample=f"{threads} chains (one per host core) x {sweeps} full sweeps of the same lattice/beta, "
                       f"C++ restatement of the reference ({'-march=native' if native else 'x86-64-v3'}); "
                       f"1 chain on 1 core: {1.0 / one:.3f} sweeps/s", seconds=dt)


def cpu_ising_baseline(w, target_seconds=10.0, threads=None):
    from oracle import oracle as orc
    L = orc.lib()
    threads = 1
    Ls = min(w["L"], 1024)
    s = np.empty(Ls * Ls, dtype=np.int8)
    L.ising_random_spins(s, Ls, SEED, 0)
    t4, t8 = C.c_uint32(), C.c_uint32()
    L.ising_thresholds(w["beta"], C.byref(t4), C.byref(t8))
    t0 = time.perf_counter()
    L.ising_checkerboard_sweeps(s, Ls, t4.value, t8.value, SEED, 0, 0, 1)
    one = time.perf_counter() - t0
    n = max(1, min(200, int(target_seconds / max(one, 1e-4))))
    t0 = time.perf_counter()
    L.ising_checkerboard_sweeps(s, Ls, t4.value, t8.value, SEED, 0, 1, n)
    dt = time.perf_counter() - t0
    return dict(value=Ls * Ls * n / dt / 1e9, unit="spin-flip attempts/ns", cores=threads, kind="port",
                sample=f"1 replica L={Ls}, {n} checkerboard sweeps on 1 core (scalar oracle)", seconds=dt)


# --------------------------------------------------------------------------- reference arm
def run_reference(args, w, rank):
    if rank != 0:
        return
    if w["kind"] != "dqmc":
        b = cpu_ising_baseline(w)
        line = dict(impl="reference", metric="Ising spin-flip attempts/ns", value=b["value"], unit=b["unit"],
                    n_gpus=args.gpus, steps=args.steps, warmup=args.warmup, ms_per_step=None,
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="u64 bit-planes",
                    data="synthetic", config=config_for(args.workload, w, args.gpus),
                    cpu_baseline={k: b[k] for k in ("value", "unit", "cores", "kind", "sample")},
                    e2e=dict(value=b["value"], unit=b["unit"], h2d_bytes_per_step=0, d2h_bytes_per_step=0))
        print(json.dumps(line))
        return
    from oracle import oracle as orc
    import montecarlo_jl_b200 as mcb
    threads = os.cpu_count() or 1
    model = mcb.HubbardModelAttractive(L=w["L"], U=w["U"])
    hop = model.hopping_exponentials(w["dtau"])
    M, N = int(round(w["beta"] / w["dtau"])), w["L"] ** 2
    try:
        orc.lib(native=True); native = True
    except Exception:
        native = False
    chains = [orc.OracleDQMC(N, M, w["safe_mult"], w["dtau"], w["U"], hop, model.rand_conf(M, SEED, c), seed=SEED,
                             chain=c, native=native) for c in range(threads)]
    for c in chains:
        c.build_stack()
    # A step = `upd` update() calls of `threads` chains (one per core): a full sweep (2M updates) where that takes
    # a few seconds at most, else a bounded part of one; chain-sweeps are counted in proportion. Exactly W warm-up
    # steps and K timed steps as asked.
    t0 = time.perf_counter()
    orc.update_many(chains, 2, threads)
    t_upd = (time.perf_counter() - t0) / 2
    budget = 120.0 / max(1, args.steps + args.warmup)            # seconds per step
    upd = int(max(1, min(2 * M, budget / max(t_upd, 1e-6))))
    for _ in range(args.warmup):
        orc.update_many(chains, upd, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.update_many(chains, upd, threads)
    dt = time.perf_counter() - t0
    value = threads * args.steps * (upd / (2.0 * M)) / dt
    sample = (f"each step = {upd} of the {2 * M} update() calls of a sweep, for {threads} chains (one per host core; "
              f"the reference is single-threaded per chain), chain-sweeps counted in proportion; C++ restatement of "
              f"MonteCarlo.jl ({'-march=native' if native else 'x86-64-v3'}), not Julia")
    line = dict(impl="reference", metric="DQMC chain-sweeps/sec", value=value, unit="chain-sweeps/s",
                n_gpus=args.gpus, steps=args.steps, warmup=args.warmup, ms_per_step=dt / args.steps * 1e3,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64", data="synthetic",
                config=config_for(args.workload, w, args.gpus),
                cpu_baseline=dict(value=value, unit="chain-sweeps/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=value, unit="chain-sweeps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


# --------------------------------------------------------------------------- our arm: DQMC
def run_dqmc(args, w, rank, local_rank, world):
    import torch
    import montecarlo_jl_b200 as mcb
    from montecarlo_jl_b200 import _lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    C_ = w["chains"]
    model = mcb.HubbardModelAttractive(L=w["L"], U=w["U"])
    mc = mcb.DQMC(model, n_chains=C_, seed=SEED, chain_offset=rank * C_, device=local_rank, beta=w["beta"],
                  delta_tau=w["dtau"], safe_mult=w["safe_mult"], check_propagation_error=False)
    N, M = mc.N, mc.M
    L = mc.L
    mc.build_stack()
    for _ in range(args.warmup):
        mc.sweep(1)
    # ---- timed region: K sweeps, device-resident state, CUDA events on the library's stream
    nl = len(_lib.LAUNCH_CLASSES)
    _lib.check(L.mcb_dqmc_profile(mc.h, 0, 1))
    st0 = mc.stats()
    with ClockSampler(local_rank) as clk:
        barrier()
        t0 = time.perf_counter()
        mc.sweep(args.steps)
        ms = L.mcb_dqmc_last_elapsed_ms(mc.h)
        barrier()
        wall = time.perf_counter() - t0
    ms = max_over_ranks(ms)
    st1 = mc.stats()
    launches = np.zeros(nl, dtype=np.int64); pms = np.zeros(nl); timed = np.zeros(nl, dtype=np.int64); flops = np.zeros(nl)
    _lib.check(L.mcb_dqmc_get_profile(mc.h, launches, pms, timed, flops))
    gpu_launches = int(launches.sum())
    acc = float((st1["acc_local"] - st0["acc_local"]).sum()) / max(1.0, float((st1["prop_local"] - st0["prop_local"]).sum()))
    value = world * C_ * args.steps / (ms * 1e-3)

    # ---- per-kernel-class durations: a second pass of the same sweeps with every launch bracketed
    prof_steps = min(args.steps, 3)
    _lib.check(L.mcb_dqmc_profile(mc.h, 1, 1))
    st2 = mc.stats()
    mc.sweep(prof_steps)
    st3 = mc.stats()
    _lib.check(L.mcb_dqmc_get_profile(mc.h, launches, pms, timed, flops))
    _lib.check(L.mcb_dqmc_profile(mc.h, 0, 1))
    acc_flop = 2.0 * N * N * float((st3["acc_local"] - st2["acc_local"]).sum())  # rank-1 updates actually applied
    shares = {}
    tot = float(pms.sum())
    for i, name in enumerate(_lib.LAUNCH_CLASSES):
        if timed[i]:
            f = flops[i] + (acc_flop if name in ("sweep_spatial", "fused") else 0.0)
            shares[name] = dict(launches_per_step=int(launches[i] // prof_steps), ms_per_step=pms[i] / prof_steps,
                                share=pms[i] / tot, avg_us=1e3 * pms[i] / timed[i],
                                tflops=f / (pms[i] * 1e-3) / 1e12 if pms[i] > 0 else None)
    top = max(shares, key=lambda k: shares[k]["share"])

    # ---- e2e: through the C ABI with pinned HOST buffers
    conf_host = torch.empty(C_ * M * N, dtype=torch.int8).pin_memory()
    g_host = torch.empty(C_ * N * N, dtype=torch.float64).pin_memory()
    conf_np, g_np = conf_host.numpy(), g_host.numpy()
    _lib.check(L.mcb_dqmc_get_conf(mc.h, conf_np))
    e2e_steps = max(3, min(args.steps, 10))

    e2e_calls = {}   # slowest wall time of each call of the step, ms (diagnostic: host-side stalls show up here)

    def timed_call(name, rc_of):
        t = time.perf_counter()
        _lib.check(rc_of())
        e2e_calls[name] = max(e2e_calls.get(name, 0.0), 1e3 * (time.perf_counter() - t))

    def e2e_step():
        timed_call("set_conf", lambda: L.mcb_dqmc_set_conf(mc.h, conf_np))        # H2D: this step's HS field
        timed_call("build_stack", lambda: L.mcb_dqmc_build_stack(mc.h))           # run!'s prologue (DQMC.jl:412-414)
        timed_call("sweep", lambda: L.mcb_dqmc_sweep(mc.h, 1, 0))                 # the sweep
        timed_call("get_greens", lambda: L.mcb_dqmc_get_greens(mc.h, g_np, 0))    # D2H: mc.s.greens
        timed_call("get_conf", lambda: L.mcb_dqmc_get_conf(mc.h, conf_np))        # D2H: mc.conf
    e2e_step()
    e2e_calls.clear()
    barrier()
    t0 = time.perf_counter()
    e2e_step_ms = []
    for _ in range(e2e_steps):
        ts = time.perf_counter()
        e2e_step()
        e2e_step_ms.append(1e3 * (time.perf_counter() - ts))
    barrier()
    e2e_dt = max_over_ranks(time.perf_counter() - t0)
    e2e_value = world * C_ * e2e_steps / e2e_dt

    hbm_peak, hbm_src = measured_peaks()
    fp64_peak = fp64_peak_tflops(torch, dev)
    F_sweep = M * float(N) ** 3 * (12.0 + 4.0 * acc + 40.0 / w["safe_mult"])
    line = {
        "metric": "DQMC chain-sweeps/sec", "value": value, "unit": "chain-sweeps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_for(args.workload, w, world), "acceptance": acc,
        "gpu_launches": gpu_launches,
        "e2e": {"value": e2e_value, "unit": "chain-sweeps/s", "h2d_bytes_per_step": int(conf_np.nbytes),
                "d2h_bytes_per_step": int(conf_np.nbytes + g_np.nbytes), "steps": e2e_steps,
                "step_ms": [round(x, 2) for x in e2e_step_ms], "slowest_call_ms": {k: round(v, 2) for k, v in e2e_calls.items()},
                "what": "set_conf(H2D) + build_stack + sweep(1) + get_greens + get_conf (D2H), wall clock, pinned host buffers"},
        "clocks": clk.summary(),
        "whole_sweep_fp64": {"algorithmic_flop_per_chain_sweep": F_sweep,
                             "achieved_tflops": F_sweep * value / 1e12,
                             "peak_tflops": fp64_peak, "frac": F_sweep * value / 1e12 / fp64_peak / world,
                             "peak_source": "cuBLAS DGEMM 4096^3 via torch.matmul, best of 5, measured in this run"},
        "kernels": shares,
        "wall_s_timed_region": wall,
    }
    k = shares[top]
    # every DQMC launch class is FP64 contraction work (wrap / stack products, QR, the delayed-update flush
    # G -= U * Rc of sweep_spatial: 2 N^2 flop per accepted flip): tensor (DMMA) roofline. The site loop inside
    # sweep_spatial is latency-bound; its fraction of the tensor peak says how far the kernel is from pipe-bound.
    line["roofline"] = {"kernel": top, "bound": "tensor", "achieved": k["tflops"], "peak": fp64_peak,
                        "unit": "TFLOP/s", "frac": (k["tflops"] or 0.0) / fp64_peak,
                        "traffic": traffic_for(args.workload, top), "avg_launch_us": k["avg_us"],
                        "share_of_step": k["share"],
                        "peak_source": "FP64 DGEMM measured in this run (MEASURED_PEAKS.json has no FP64 entry)"}
    # the one collective of the path: all-reduce of the observable moments (sum x, sum x^2 of <G>,
    # occupation, CDC, SDC, pairing, energies per chain, plus the count) over the ranks, once per run after the sweeps.
    # Outside the timed region of `value`; its own time is reported here (NCCL when world > 1).
    from montecarlo_jl_b200 import sharding
    mc.enable_measurements()
    mc.measure()
    per_chain = {k2: v for k2, v in mc.observables().items() if k2 != "count"}
    rate = (st1["acc_local"] - st0["acc_local"]) / np.maximum(1, st1["prop_local"] - st0["prop_local"])
    per_chain["acceptance"] = rate
    mom, _ = sharding.local_moments(per_chain)
    payload = int(sum(v.size for v in mom.values()) + 1) * 8
    barrier()
    t0 = time.perf_counter()
    red, n_red = sharding.reduce_observables(per_chain, device=dev)
    barrier()
    red_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    line["observable_reduction"] = {"chains": n_red, "payload_bytes": payload, "ms": red_ms,
                                    "acceptance_mean": float(red["acceptance"][0]),
                                    "acceptance_stderr": float(red["acceptance"][1]),
                                    "occupation_mean": float(red["n"][0]),
                                    "collective": "nccl all_reduce(sum) of moments" if world > 1 else "none (1 rank)"}
    # the other BASELINE.json configs, so that the driver's records (N = 1, 2, 4, 8) carry them too
    if args.workload == "dqmc_A512" and not args.no_extra_configs:
        line["configs"] = {
            "dqmc_C128": quick_dqmc("dqmc_C128", rank, local_rank, world, max_over_ranks, barrier, fp64_peak, steps=8, warmup=4),
            "dqmc_D32": quick_dqmc("dqmc_D32", rank, local_rank, world, max_over_ranks, barrier, fp64_peak, steps=2),
            "ising64": quick_ising("ising64", rank, local_rank, world, max_over_ranks, barrier, steps=5),
        }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        b = cpu_dqmc_baseline(w)
        line["cpu_baseline"] = {k2: b[k2] for k2 in ("value", "unit", "cores", "kind", "sample")}
    mc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line))


# --------------------------------------------------------------------------- the other BASELINE configs, short runs
def quick_dqmc(name, rank, local_rank, world, max_over_ranks, barrier, fp64_peak, steps, warmup=3):
    """Short device-timed run of another DQMC workload for the `configs` object of the default line: same
    timing rules as the main number (W >= 3 warm-up sweeps, CUDA events on the library's stream, max over ranks)."""
    import montecarlo_jl_b200 as mcb
    from montecarlo_jl_b200 import _lib
    w = WORKLOADS[name]
    C_ = w["chains"]
    model = mcb.HubbardModelAttractive(L=w["L"], U=w["U"])
    mc = mcb.DQMC(model, n_chains=C_, seed=SEED, chain_offset=rank * C_, device=local_rank, beta=w["beta"],
                  delta_tau=w["dtau"], safe_mult=w["safe_mult"], check_propagation_error=False)
    L, N, M = mc.L, mc.N, mc.M
    mc.build_stack()
    mc.sweep(warmup)
    st0 = mc.stats()
    barrier()
    mc.sweep(steps)
    ms = max_over_ranks(L.mcb_dqmc_last_elapsed_ms(mc.h))
    barrier()
    st1 = mc.stats()
    acc = float((st1["acc_local"] - st0["acc_local"]).sum()) / max(1.0, float((st1["prop_local"] - st0["prop_local"]).sum()))
    value = world * C_ * steps / (ms * 1e-3)
    nl = len(_lib.LAUNCH_CLASSES)
    launches = np.zeros(nl, dtype=np.int64); pms = np.zeros(nl); timed = np.zeros(nl, dtype=np.int64); flops = np.zeros(nl)
    _lib.check(L.mcb_dqmc_profile(mc.h, 1, 1))
    st2 = mc.stats()
    mc.sweep(1)
    st3 = mc.stats()
    _lib.check(L.mcb_dqmc_get_profile(mc.h, launches, pms, timed, flops))
    _lib.check(L.mcb_dqmc_profile(mc.h, 0, 1))
    acc_flop = 2.0 * N * N * float((st3["acc_local"] - st2["acc_local"]).sum())
    kern = {}
    for i, cls in enumerate(_lib.LAUNCH_CLASSES):
        if timed[i] and pms[i] > 0:
            f = flops[i] + (acc_flop if cls in ("sweep_spatial", "fused") else 0.0)
            kern[cls] = dict(ms_per_step=float(pms[i]), share=float(pms[i] / pms.sum()), launches_per_step=int(launches[i]),
                             tflops=f / (pms[i] * 1e-3) / 1e12, frac=f / (pms[i] * 1e-3) / 1e12 / fp64_peak)
    top = max(kern, key=lambda c: kern[c]["share"])
    F_sweep = M * float(N) ** 3 * (12.0 + 4.0 * acc + 40.0 / w["safe_mult"])
    mc.close()
    return {"metric": "DQMC chain-sweeps/sec", "value": value, "unit": "chain-sweeps/s", "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "config": config_for(name, w, world), "acceptance": acc,
            "whole_sweep_frac": F_sweep * value / 1e12 / fp64_peak / world, "dominant_kernel": top, "kernels": kern}


def quick_ising(name, rank, local_rank, world, max_over_ranks, barrier, steps, warmup=3):
    import montecarlo_jl_b200 as mcb
    w = WORKLOADS[name]
    R, Ls = w["replicas"], w["L"]
    sim = mcb.MC(mcb.IsingModel(Ls), beta=w["beta"], n_replicas=R, seed=SEED, replica_offset=rank * R, device=local_rank)
    for _ in range(warmup):
        sim.sweep(10)
    barrier()
    sim.sweep(steps * 10)
    ms = max_over_ranks(sim.L.mcb_ising_last_elapsed_ms(sim.h))
    barrier()
    sim.close()
    flips = float(world) * R * Ls * Ls * steps * 10
    hbm_peak, _ = measured_peaks()
    gbs = flips / world * (3.0 / 8.0) / (ms * 1e-3) / 1e9
    return {"metric": "Ising spin-flip attempts/ns", "value": flips / (ms * 1e-3) / 1e9, "unit": "spin-flip attempts/ns",
            "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "config": config_for(name, w, world),
            "dominant_kernel": "ising_halfsweep_kernel", "hbm_frac_algorithmic": gbs / hbm_peak}


# --------------------------------------------------------------------------- our arm: Ising
def run_ising(args, w, rank, local_rank, world):
    import torch
    import montecarlo_jl_b200 as mcb
    from montecarlo_jl_b200 import _lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    R, Ls = w["replicas"], w["L"]
    sim = mcb.MC(mcb.IsingModel(Ls), beta=w["beta"], n_replicas=R, seed=SEED, replica_offset=rank * R,
                 device=local_rank)
    L = sim.L
    sweeps_per_step = 10
    for _ in range(args.warmup):
        sim.sweep(sweeps_per_step)
    l0 = L.mcb_ising_launch_count(sim.h)
    with ClockSampler(local_rank) as clk:
        barrier()
        sim.sweep(args.steps * sweeps_per_step)
        ms = L.mcb_ising_last_elapsed_ms(sim.h)
        barrier()
    ms = max_over_ranks(ms)
    gpu_launches = int(L.mcb_ising_launch_count(sim.h) - l0)
    flips = float(world) * R * Ls * Ls * args.steps * sweeps_per_step
    value = flips / (ms * 1e-3) / 1e9
    # e2e: spins up, sweeps, spins + (E, M) down, in the bit-packed wire format (1 bit per spin over PCIe)
    nch = L.mcb_ising_spin_chunks(sim.h)
    host = torch.empty(R * nch, dtype=torch.int64).pin_memory()
    hnp = host.numpy().view(np.uint64)
    _lib.check(L.mcb_ising_get_spins_bits(sim.h, hnp))
    E = np.empty(R, dtype=np.int64); Msum = np.empty(R, dtype=np.int64)
    e2e_steps = max(3, min(args.steps, 10))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        _lib.check(L.mcb_ising_set_spins_bits(sim.h, hnp))
        _lib.check(L.mcb_ising_sweep(sim.h, sweeps_per_step))
        _lib.check(L.mcb_ising_measure(sim.h, E, Msum))
        _lib.check(L.mcb_ising_get_spins_bits(sim.h, hnp))
    barrier()
    dt = max_over_ranks(time.perf_counter() - t0)
    hbm_peak, hbm_src = measured_peaks()
    gbs = flips / world * (3.0 / 8.0) / (ms * 1e-3) / 1e9   # 3/8 byte per spin per sweep
    line = {"metric": "Ising spin-flip attempts/ns", "value": value, "unit": "spin-flip attempts/ns", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64 bit-planes", "data": "synthetic",
            "config": config_for(args.workload, w, world),
            "gpu_launches": gpu_launches,
            "e2e": {"value": world * R * Ls * Ls * e2e_steps * sweeps_per_step / dt / 1e9, "unit": "spin-flip attempts/ns",
                    "h2d_bytes_per_step": int(hnp.nbytes), "d2h_bytes_per_step": int(hnp.nbytes + 16 * R)},
            "clocks": clk.summary(),
            "roofline": {"kernel": "ising_halfsweep_kernel", "bound": "hbm", "achieved": gbs, "peak": hbm_peak,
                         "unit": "GB/s", "frac": gbs / hbm_peak, "traffic": traffic_for(args.workload, "ising"),
                         "peak_source": hbm_src,
                         "avg_launch_us": 1e3 * ms / max(1, gpu_launches),
                         # what actually bounds the kernel (ncu, profiles/r02u_ncu_full_summary.txt): the planes stay in L2,
                         # the integer ALU pipe is 71 % busy (24.7 thread instructions per spin, two thirds of them the ten
                         # Philox4x32 rounds shared by four spins); ALU ceiling = value / 0.71
                         "bound_detail": {"limiter": "integer ALU pipe", "alu_pipe_pct_of_peak": 71.1,
                                          "thread_instructions_per_spin": 24.7, "philox_rounds_per_spin": 2.5,
                                          "alu_ceiling_flips_per_ns": value / world / 0.711 * world,
                                          "source": "profiles/r02u_ncu_full_summary.txt"}}}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        b = cpu_ising_baseline(w)
        line["cpu_baseline"] = {k2: b[k2] for k2 in ("value", "unit", "cores", "kind", "sample")}
    sim.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="dqmc_A512", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-configs", action="store_true",
                    help="skip the short runs of the other BASELINE configs that the default line carries in `configs`")
    args = ap.parse_args()
    rank, local_rank, world = dist_env()
    w = WORKLOADS[args.workload]
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args, w, rank)
        return
    if w["kind"] == "dqmc":
        run_dqmc(args, w, rank, local_rank, world)
    else:
        run_ising(args, w, rank, local_rank, world)


if __name__ == "__main__":
    main()
