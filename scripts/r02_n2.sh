#!/bin/bash
# 2-GPU box: data-parallel correctness (tests) + bench at N = 2 with the D x h gradients as bf16 / fp32 on the wire
TAG=${1:-r02n2}
O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_host_main.py -q -rs -k "data_parallel or two_gpus" ) > $O/${TAG}_pytest.log 2>&1; tail -12 $O/${TAG}_pytest.log
# fp32 tolerance on the tensor cores ("dtype" 2): the new parity cases, then a bench line (one GPU)
( CUDA_VISIBLE_DEVICES=0 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py -q -rs -k "f32x3" ) > $O/${TAG}_pytest_f32x3.log 2>&1; tail -15 $O/${TAG}_pytest_f32x3.log
( CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --dtype f32x3 --cells 65536 --steps 10 --warmup 3 --no-cpu-baseline ) > $O/${TAG}_bench_f32x3.log 2>&1; tail -c 1500 $O/${TAG}_bench_f32x3.log
for W in 1 0; do
  MMV_COMM_BF16=$W timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus 2 --steps 60 --warmup 5 --no-cpu-baseline > $O/${TAG}_bench_wire$W.log 2>&1
  grep -a '^{' $O/${TAG}_bench_wire$W.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); k=j["kernels"]; print("comm_bf16='$W' N=%d %.4f ms/step %.0f cells/s e2e %.0f dp_check %s allreduce %.3f senc_bwd %.3f adam %.3f" % (j["n_gpus"], j["ms_per_step"], j["value"], j["e2e"]["value"], j.get("dp_check"), k["allreduce"]["ms_per_step"], k["senc_bwd"]["ms_per_step"], k["clip_adam"]["ms_per_step"]))'
done 2>&1 | tee $O/${TAG}_summary.txt
# (one GPU) programmatic dependent launch on / off, K4 excluded from it
for rep in 1 2; do for P in 0 1; do
  ( CUDA_VISIBLE_DEVICES=0 MMV_PDL=$P timeout 300 python bench.py --cells 131072 --steps 40 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_pdl${P}_${rep}.log 2>&1
  grep -a '^{' $O/${TAG}_pdl${P}_${rep}.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); k=j["kernels"]; print("MMV_PDL='$P' rep '$rep': step %.4f e2e %.4f nb_likelihood %.4f launches %.0f" % (j["ms_per_step"], j["e2e"]["ms_per_step"], k["nb_likelihood"]["ms_per_step"], j["gpu_launches"]/j["steps"]))'
done; done 2>&1 | tee $O/${TAG}_pdl_ab.txt
