#!/bin/bash
# A/B of one environment knob on one box with the in-tree library: per-kernel times of the default bench
# usage: r02_env_ab.sh <tag> <VAR> <value...>      (each value twice, interleaved)
TAG=$1; VAR=$2; shift 2
O=gpurun_out; mkdir -p $O
for rep in 1 2; do
for V in "$@"; do
  ( env $VAR=$V timeout 300 python bench.py --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_${V}_${rep}.log 2>&1
  python - <<PY | tee -a $O/${TAG}.txt
import json
try:
    d=json.loads([l for l in open('$O/${TAG}_${V}_${rep}.log') if l.startswith('{')][-1])
    k=d['kernels']
    print('$VAR=$V rep $rep: step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], {n: k[n]['ms_per_step'] for n in ('panel_build','senc_fwd','senc_bwd','nb_likelihood','colsum','gemm_other') if n in k})
except Exception as e:
    print('$VAR=$V: bench failed', e)
PY
done
done
