#!/bin/bash
# one-box check of programmatic dependent launch: full GPU tests with PDL on (the default), then the default bench with
# MMV_PDL=0 / 1 alternating (same library, same box)
TAG=${1:-r02d}
O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -rs ) > $O/${TAG}_pytest.log 2>&1; tail -8 $O/${TAG}_pytest.log
( time timeout 400 python __graft_entry__.py --smoke ) > $O/${TAG}_smoke.log 2>&1; tail -5 $O/${TAG}_smoke.log
for rep in 1 2; do
for P in 0 1; do
  ( MMV_PDL=$P timeout 300 python bench.py --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_pdl${P}_${rep}.log 2>&1
  python - <<PY
import json
try:
    d=json.loads([l for l in open('$O/${TAG}_pdl${P}_${rep}.log') if l.startswith('{')][-1])
    k=d['kernels']
    print('MMV_PDL=$P rep $rep: step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], {n: k[n]['ms_per_step'] for n in ('nb_likelihood','gemm_dec_last_fwd','senc_fwd','senc_bwd','gemm_other','clip_adam','panel_build') if n in k}, 'launches', d['gpu_launches']/d['steps'])
except Exception as e:
    print('MMV_PDL=$P: bench failed', e)
PY
done
done 2>&1 | tee $O/${TAG}_pdl_ab.txt
( MMV_PDL=1 timeout 300 python bench.py --model vmf --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_vmf_pdl1.log 2>&1; tail -c 300 $O/${TAG}_vmf_pdl1.log
( MMV_PDL=0 timeout 300 python bench.py --model vmf --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_vmf_pdl0.log 2>&1; tail -c 300 $O/${TAG}_vmf_pdl0.log
