#!/bin/bash
mkdir -p gpurun_out
for v in 4 8; do
echo "== step warps $v"
MCB_QRB_PANEL_WARPS=$v timeout 600 python -m pytest tests -m gpu -x -q -k "udt or noalloc or config_D" 2>&1 | tail -1
MCB_QRB_PANEL_WARPS=$v timeout 300 python - 2>&1 <<'PY' | grep qrb_panel | tail -1
import sys
sys.path.insert(0, '.')
from montecarlo_jl_b200 import _lib
_lib.LIB_PATH = 'profiles/microbench/libmcb200_timing.so'
sys.argv = ['x', '400', '32']
exec(open('profiles/microbench/udt400.py').read())
PY
MCB_QRB_PANEL_WARPS=$v timeout 300 python bench.py --workload dqmc_D32 --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['config']['workload'], d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],1) for k,v in d['kernels'].items()})"
done
