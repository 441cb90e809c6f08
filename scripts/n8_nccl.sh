#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
TAG=default NCCL_DEBUG=INFO $TR scripts/ar_probe.py > gpurun_out/arp_default.log 2>&1
grep -a "ar_probe" gpurun_out/arp_default.log
grep -a -i "nvls\|channels\|Connected" gpurun_out/arp_default.log | grep -a "\[0\]" | head -12
TAG=ring NCCL_ALGO=Ring $TR scripts/ar_probe.py 2>&1 | grep -a "ar_probe"
TAG=cta16 NCCL_MAX_CTAS=16 $TR scripts/ar_probe.py 2>&1 | grep -a "ar_probe"
for v in "NCCL_ALGO=Ring" "NCCL_MAX_CTAS=16"; do
  env $v MMV_COMM_TRACE=20 $TR bench.py --gpus 8 --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/n8_$v.log 2>&1
  echo "== $v: $(grep -a '^{' gpurun_out/n8_$v.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); print(j["ms_per_step"], j["value"])')"
  grep -a "trace\]" gpurun_out/n8_$v.log | grep -a "comm\|first-layer backward\|joined\|step_end"
done
