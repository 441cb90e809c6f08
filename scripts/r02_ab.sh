#!/bin/bash
# A/B of library builds under gpurun_libs/*.so on one box: per-kernel times of the default bench (usage: r02_ab.sh <tag> <names...>)
TAG=$1; shift
O=gpurun_out; mkdir -p $O
cp mm-vae_b200/libmmvae_b200.so /tmp/lib_orig.so
for rep in 1 2; do
for V in "$@"; do
  cp gpurun_libs/$V.so mm-vae_b200/libmmvae_b200.so
  ( timeout 300 python bench.py --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_${V}_${rep}.log 2>&1
  python - <<PY
import json
try:
    d=json.loads([l for l in open('$O/${TAG}_${V}_${rep}.log') if l.startswith('{')][-1])
    k=d['kernels']
    print('$V rep $rep: step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], {n: k[n]['ms_per_step'] for n in ('nb_likelihood','gemm_dec_last_dx','senc_fwd','senc_bwd','panel_build','colsum_dlogits','gemm_other') if n in k}, 'launches', d['gpu_launches']/d['steps'])
except Exception as e:
    print('$V: bench failed', e)
PY
done
done
cp /tmp/lib_orig.so mm-vae_b200/libmmvae_b200.so
