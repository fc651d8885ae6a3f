#!/bin/bash
mkdir -p gpurun_out
for v in 1 0 1 0; do echo "== own pool $v"; MCB_SCRATCH_POOL=$v timeout 300 python profiles/microbench/e2e_breakdown.py dqmc_C128 24 2>&1 | awk '{print $10}' | sort -n | tr '\n' ' '; echo; done 2>&1 | tee gpurun_out/exp_e2e.txt
