#!/bin/bash
# last pass of a round on a B200: the whole GPU suite, smoke(), the default bench line and the lines of configs C / D
tag=${1:-final}
o=gpurun_out
mkdir -p $o
timeout 1200 python -m pytest tests -m gpu -q > $o/${tag}_pytest_gpu.log 2>&1; tail -2 $o/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py 2> $o/${tag}_bench_A.err | tail -1 > $o/${tag}_bench_dqmc_A512.json
for w in dqmc_C128 dqmc_D32; do
  timeout 600 python bench.py --workload $w --no-extra-configs 2> $o/${tag}_bench_$w.err | tail -1 > $o/${tag}_bench_$w.json
done
python - <<PY
import json
for w in ("dqmc_A512", "dqmc_C128", "dqmc_D32"):
    d = json.loads(open("$o/${tag}_bench_%s.json" % w).read())
    print(w, round(d["value"], 1), round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["whole_sweep_fp64"]["frac"], 4),
          {k: round(v["ms_per_step"], 1) for k, v in d["kernels"].items()}, {k: round(v["value"], 1) for k, v in d.get("configs", {}).items()})
PY
