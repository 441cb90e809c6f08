#!/bin/bash
# 8-GPU box, fourth run: two parts (the new default) vs one, and NCCL's NVLS algorithm instead of the ring
TAG=${1:-r02n8d}
O=gpurun_out; mkdir -p $O
run() {  # name, env, extra bench args
  env $2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus 8 --steps 60 --warmup 5 --no-cpu-baseline $3 > $O/${TAG}_$1.log 2>&1
  grep -a "trace\]" $O/${TAG}_$1.log > $O/${TAG}_$1_trace.txt
  grep -a '^{' $O/${TAG}_$1.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); k=j["kernels"]; print("'$1': N=%d %.4f ms/step %.0f cells/s e2e %.0f (%.4f ms) secondary %s dp_check %s allreduce %.3f senc_bwd %.3f adam %.3f" % (j["n_gpus"], j["ms_per_step"], j["value"], j["e2e"]["value"], j["e2e"]["ms_per_step"], j.get("secondary",{}).get("value"), j.get("dp_check",{}).get("worst_grad_rel_err"), k["allreduce"]["ms_per_step"], k["senc_bwd"]["ms_per_step"], k["clip_adam"]["ms_per_step"]))'
}
{
run default "MMV_COMM_TRACE=30" ""
run parts1 "MMV_DW1_PARTS=1 MMV_COMM_TRACE=30" ""
run nvls "NCCL_ALGO=NVLS MMV_COMM_TRACE=30" ""
run cfg5 "" "--config 5"
} 2>&1 | tee $O/${TAG}_summary.txt
