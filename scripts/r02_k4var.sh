#!/bin/bash
# K4 block-size variants: parity at D = 30 000 + per-kernel time (usage: r02_k4var.sh <tag> <variants...>)
TAG=$1; shift
O=gpurun_out
mkdir -p $O
for V in "$@"; do
  export MMV_K4_VARIANT=$V
  ( timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py -q -x -k "nb or full_size_minibatch or properties or fused or streamed" ) > $O/${TAG}_v${V}_pytest.log 2>&1; tail -3 $O/${TAG}_v${V}_pytest.log
  ( timeout 300 python bench.py --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_v${V}_bench.log 2>&1
  python - <<PY
import json
try:
    d=json.loads([l for l in open('$O/${TAG}_v${V}_bench.log') if l.startswith('{')][-1])
    print('variant $V: step', d['ms_per_step'], 'nb_likelihood', d['kernels']['nb_likelihood']['ms_per_step'], d['kernels']['nb_likelihood']['frac'], 'fwd gemm', d['kernels']['gemm_dec_last_fwd']['ms_per_step'])
except Exception as e:
    print('variant $V: bench failed', e)
PY
done
