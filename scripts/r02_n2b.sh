#!/bin/bash
# 2-GPU box, second run: the data-parallel tests on the final code (clip-norm partials from the bf16 unpack pass), the 2-GPU
# main, bench at N = 2; then (one GPU) the panel_build L2-prefetch A/B
TAG=${1:-r02n2b}
O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_host_main.py -q -rs -k "data_parallel or two_gpus" ) > $O/${TAG}_pytest.log 2>&1; tail -6 $O/${TAG}_pytest.log
for W in 1 0; do
  MMV_COMM_BF16=$W timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus 2 --steps 60 --warmup 5 --no-cpu-baseline > $O/${TAG}_bench_wire$W.log 2>&1
  grep -a '^{' $O/${TAG}_bench_wire$W.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); k=j["kernels"]; print("comm_bf16='$W' N=%d %.4f ms/step %.0f cells/s e2e %.0f dp_check %s allreduce %.3f senc_bwd %.3f adam %.3f" % (j["n_gpus"], j["ms_per_step"], j["value"], j["e2e"]["value"], j.get("dp_check",{}).get("worst_grad_rel_err"), k["allreduce"]["ms_per_step"], k["senc_bwd"]["ms_per_step"], k["clip_adam"]["ms_per_step"]))'
done 2>&1 | tee $O/${TAG}_summary.txt
CUDA_VISIBLE_DEVICES=0 bash scripts/r02_ab.sh ${TAG}_ab base pf 2>&1 | tee $O/${TAG}_panel_prefetch_ab.txt
