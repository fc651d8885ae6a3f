#!/bin/bash
# the default bench line on all 8 GPUs of a box, launched the way the driver does (configs C / D / Ising ride along)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 2> gpurun_out/r05_8gpu.err | tail -1 > gpurun_out/r05_bench_A_8gpu.json
python -c "
import json
d=json.loads(open('gpurun_out/r05_bench_A_8gpu.json').read())
print('A 8gpu', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {k:round(v['value'],1) for k,v in d.get('configs',{}).items()}, d.get('observable_reduction',{}).get('collective'))
"
