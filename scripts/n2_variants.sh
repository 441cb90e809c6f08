#!/bin/bash
# N-GPU bench under a few settings (usage: n2_variants.sh <ngpus> <tag> VAR=val ...): one traced run per setting
N=${1:-2}; TAG=${2:-v}; shift 2
run() {
  name=$1
  env "$1" MMV_COMM_TRACE=20 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_${name}.log 2>&1
  echo "== $name: $(grep -a '^{' gpurun_out/${TAG}_${name}.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); print(j["ms_per_step"], j["value"], j["e2e"]["value"])')"
  grep -a "trace\]" gpurun_out/${TAG}_${name}.log | grep -a "comm\|first-layer backward\|joined\|step_end\|decoder-last\|gradient buffer"
  grep -a -i "error\|warn" gpurun_out/${TAG}_${name}.log | head -5
}
for v in "$@"; do run "$v"; done
