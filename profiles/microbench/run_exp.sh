#!/bin/bash
# scratch driver of one gpurun call (experiments); outputs under gpurun_out/
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "udt or config or rdivp or calculate_greens or noalloc" > gpurun_out/exp_tests.txt 2>&1
tail -3 gpurun_out/exp_tests.txt
timeout 300 python - > gpurun_out/exp_udt_timing.txt 2>&1 <<'PY'
import sys
sys.path.insert(0, '.')
from montecarlo_jl_b200 import _lib
_lib.LIB_PATH = 'profiles/microbench/libmcb200_timing.so'
sys.argv = ['x', '400', '32']
exec(open('profiles/microbench/udt400.py').read())
sys.argv = ['x', '144', '128']
exec(open('profiles/microbench/udt400.py').read())
PY
grep -v "^$" gpurun_out/exp_udt_timing.txt | tail -12
for w in dqmc_C128 dqmc_D32; do timeout 300 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['config']['workload'], d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],1) for k,v in d['kernels'].items()})"; done > gpurun_out/exp_bench.txt 2>&1
cat gpurun_out/exp_bench.txt
