#!/bin/bash
# usage: run_bench.sh TAG  -> tests + bench + kernel table
TAG=$1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$TAG.log 2>&1
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$TAG.log").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["last_loss"])
for k,v in d["kernels"].items(): print(k, v["ms_per_launch_group"], v["share"], v.get("frac"))
PY
