#!/bin/bash
# quick iteration check: parity tests matching $2 (default nb), then the nb bench with the kernel table (usage: gpu_quick.sh <tag> [k-expr])
TAG=${1:-q}; K=${2:-nb}
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py -x -q -k "$K" 2>&1 | tail -4
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$TAG.log 2>&1
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_$TAG.log") if l.startswith("{")][-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["last_loss"])
for k,v in d["kernels"].items(): print(k, v["ms_per_launch_group"], v["share"], v.get("frac"))
PY
