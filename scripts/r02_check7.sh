#!/bin/bash
# A/B of library builds (gpurun_libs/*.so) on one box for the nb AND vmf likelihood kernels, then the parity / edge tests
# with the in-tree library (usage: r02_check7.sh <tag> <names...>)
TAG=${1:-r02h}; shift
O=gpurun_out; mkdir -p $O
bash scripts/r02_ab.sh ${TAG}_ab "$@" 2>&1 | tee $O/${TAG}_ab.txt
cp mm-vae_b200/libmmvae_b200.so /tmp/lib_orig2.so
for V in "$@"; do
  cp gpurun_libs/$V.so mm-vae_b200/libmmvae_b200.so
  ( timeout 300 python bench.py --model vmf --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_vmf_${V}.log 2>&1
  python - <<PY | tee -a $O/${TAG}_ab.txt
import json
try:
    d=json.loads([l for l in open('$O/${TAG}_vmf_${V}.log') if l.startswith('{')][-1])
    k=d['kernels']
    print('vmf $V: step', d['ms_per_step'], {n: k[n]['ms_per_step'] for n in k if 'lik' in n})
except Exception as e:
    print('vmf $V: bench failed', e)
PY
done
cp /tmp/lib_orig2.so mm-vae_b200/libmmvae_b200.so
( time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py -m gpu -q -x ) > $O/${TAG}_pytest.log 2>&1; tail -5 $O/${TAG}_pytest.log
