#!/bin/bash
# both arms of bench.py on 2 GPUs, launched the way the driver does
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 2> gpurun_out/r04_2gpu.err | tail -1 > gpurun_out/r04_bench_A_2gpu.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>> gpurun_out/r04_2gpu.err | tail -1 > gpurun_out/r04_bench_ref_2gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 profiles/microbench/allreduce_2gpu.py > gpurun_out/r04_allreduce_2gpu.txt 2>&1
python -c "
import json
d=json.loads(open('gpurun_out/r04_bench_A_2gpu.json').read())
print('A 2gpu', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {k:round(v['value'],1) for k,v in d.get('configs',{}).items()}, d.get('observable_reduction',{}).get('collective'))
print(open('gpurun_out/r04_bench_ref_2gpu.json').read()[:300])
"
tail -3 gpurun_out/r04_allreduce_2gpu.txt
