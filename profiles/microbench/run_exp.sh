#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "udt or calculate_greens or config_C or noalloc" 2>&1 | tail -1
timeout 300 python - 2>&1 <<'PY' | grep "udt_resident cycles" | tail -1
import sys
sys.path.insert(0, '.')
from montecarlo_jl_b200 import _lib
_lib.LIB_PATH = 'profiles/microbench/libmcb200_timing.so'
sys.argv = ['x', '144', '128']
exec(open('profiles/microbench/udt400.py').read())
PY
timeout 300 python bench.py --workload dqmc_C128 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['config']['workload'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {k:round(v['ms_per_step'],2) for k,v in d['kernels'].items()})"
