#!/bin/bash
# usage: n8_trace.sh <tag>: traced bench at N=1 and N=8
TAG=${1:-t}
MMV_COMM_TRACE=20 python bench.py --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_n1.log 2>&1
grep -a "trace\]" gpurun_out/${TAG}_n1.log
grep -a '^{' gpurun_out/${TAG}_n1.log | cut -c1-200
MMV_COMM_TRACE=20 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_n8.log 2>&1
grep -a "trace\]" gpurun_out/${TAG}_n8.log
grep -a '^{' gpurun_out/${TAG}_n8.log | cut -c1-200
