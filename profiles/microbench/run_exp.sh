#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "udt or rdivp or calculate_greens or noalloc or config" > gpurun_out/exp_tests.txt 2>&1
tail -3 gpurun_out/exp_tests.txt
timeout 300 python profiles/microbench/canary_D.py 1 2>&1 | tail -2
timeout 300 python - 2>&1 <<'PY' | grep qrb_panel
import sys
sys.path.insert(0, '.')
from montecarlo_jl_b200 import _lib
_lib.LIB_PATH = 'profiles/microbench/libmcb200_timing.so'
sys.argv = ['x', '400', '32']
exec(open('profiles/microbench/udt400.py').read())
PY
timeout 300 python bench.py --workload dqmc_D32 --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['config']['workload'], d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],1) for k,v in d['kernels'].items()})"
