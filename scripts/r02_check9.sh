#!/bin/bash
# side stream A/B (MMV_SIDE=0/1) for nb and vmf on one box + the parity / edge / step tests
TAG=${1:-r02l}
O=gpurun_out; mkdir -p $O
bash scripts/r02_env_ab.sh ${TAG}_side MMV_SIDE 0 1
for V in 0 1; do
  ( MMV_SIDE=$V timeout 300 python bench.py --model vmf --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_vmf_${V}.log 2>&1
  python - <<PY | tee -a $O/${TAG}_side.txt
import json
try:
    d=json.loads([l for l in open('$O/${TAG}_vmf_${V}.log') if l.startswith('{')][-1])
    k=d['kernels']
    print('vmf MMV_SIDE=$V: step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], {n: k[n]['ms_per_step'] for n in ('vmf_likelihood','colsum_dlogits','gemm_other','colsum','panel_build','senc_bwd') if n in k})
except Exception as e:
    print('vmf $V: bench failed', e)
PY
done
( time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py -m gpu -q -x ) > $O/${TAG}_pytest.log 2>&1; tail -3 $O/${TAG}_pytest.log
