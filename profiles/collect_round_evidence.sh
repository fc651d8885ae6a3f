#!/bin/bash
# One pass on a B200: GPU test suite, bench lines of every workload, ncu launch lists and full-set captures of the
# kernels that dominate configs C / D and the Ising workload. Everything lands in gpurun_out/<tag>_*; the summaries
# under profiles/ are made from those on the CPU box (ncu_summary.py, launch_summary.py).
# usage: bash profiles/collect_round_evidence.sh r02u
tag=${1:-evidence}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests -m gpu -q > $o/${tag}_pytest_gpu.log 2>&1; tail -2 $o/${tag}_pytest_gpu.log
timeout 600 python bench.py 2> $o/${tag}_bench_A.err | tail -1 > $o/${tag}_bench_dqmc_A512.json
for w in dqmc_C128 dqmc_D32 ising64; do
  timeout 600 python bench.py --workload $w --no-extra-configs 2> $o/${tag}_bench_$w.err | tail -1 > $o/${tag}_bench_$w.json
done
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 > $o/${tag}_bench_reference.json
# launch lists (durations are cold-cache and serialised under ncu: shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $o/${tag}_launches_C.csv python profiles/microbench/sweep144.py > /dev/null 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $o/${tag}_launches_D.csv python profiles/microbench/sweep400.py > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $o/${tag}_launches_A.csv python bench.py --steps 1 --warmup 1 --no-extra-configs --no-cpu-baseline > /dev/null 2>&1
# full-set captures
timeout 300 ncu --set full --import-source on --clock-control none -k regex:segment_resident -s 1 -c 1 -f -o $o/${tag}_C_segment_resident python profiles/microbench/sweep144.py > /dev/null 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:sweep_delayed -s 2 -c 1 -f -o $o/${tag}_D_sweep_delayed python profiles/microbench/sweep400.py > /dev/null 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:gemm_tma -s 2 -c 1 -f -o $o/${tag}_D_gemm_tma python profiles/microbench/sweep400.py > /dev/null 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:qrb_panel -s 5 -c 1 -f -o $o/${tag}_D_qrb_panel python profiles/microbench/udt400.py 400 32 > /dev/null 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:qrb_formq -s 1 -c 1 -f -o $o/${tag}_D_qrb_formq python profiles/microbench/udt400.py 400 32 > /dev/null 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:segment_kernel -s 4 -c 1 -f -o $o/${tag}_A_segment_kernel python bench.py --steps 1 --warmup 1 --no-extra-configs --no-cpu-baseline > /dev/null 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:ising_halfsweep -s 4 -c 1 -f -o $o/${tag}_ising_halfsweep python bench.py --workload ising64 --steps 1 --warmup 1 --no-extra-configs --no-cpu-baseline > /dev/null 2>&1
ls -la $o | grep ${tag}_ | awk '{print $5, $9}'
