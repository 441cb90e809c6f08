#!/bin/bash
# bench at N = 2, 4, 8 (usage: scale_run.sh <tag>)
TAG=${1:-s}
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_n$N.log 2>&1
  grep -a '^{' gpurun_out/${TAG}_n$N.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); print("N=%d %.4f ms/step  %.0f cells/s  e2e %.0f" % (j["n_gpus"], j["ms_per_step"], j["value"], j["e2e"]["value"]))'
done
