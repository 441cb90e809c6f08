#!/bin/bash
# 8-GPU box, second run: comm stream at the highest priority (default) vs default priority, and the C++ main with --ngpu 8
# long enough (24 epochs) for the steady-state rate after the first epoch
TAG=${1:-r02n8b}
O=gpurun_out; mkdir -p $O
run() {  # name, env, extra bench args
  env $2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus 8 --steps 60 --warmup 5 --no-cpu-baseline $3 > $O/${TAG}_$1.log 2>&1
  grep -a "trace\]" $O/${TAG}_$1.log > $O/${TAG}_$1_trace.txt
  grep -a '^{' $O/${TAG}_$1.log | python -c 'import sys,json; j=json.loads(sys.stdin.readline()); k=j["kernels"]; print("'$1': N=%d %.4f ms/step %.0f cells/s e2e %.0f (%.4f ms) secondary %s dp_check %s allreduce %.3f senc_bwd %.3f adam %.3f numa %s" % (j["n_gpus"], j["ms_per_step"], j["value"], j["e2e"]["value"], j["e2e"]["ms_per_step"], j.get("secondary",{}).get("value"), j.get("dp_check",{}).get("worst_grad_rel_err"), k["allreduce"]["ms_per_step"], k["senc_bwd"]["ms_per_step"], k["clip_adam"]["ms_per_step"], j["config"].get("numa_bind")))'
}
{
run prio_high "MMV_COMM_TRACE=30" ""
run prio_default "MMV_COMM_PRIORITY=0 MMV_COMM_TRACE=30" ""
} 2>&1 | tee $O/${TAG}_summary.txt
( time timeout 600 mm-vae_b200/host/bin/nb_vae --synthetic 524288,30000,0.05 --out /tmp/nbvae8 --batch_size 4096 --mean_encoding 1024,256 \
    --mean_decoding 256,1024 --mean_latent 32 --max_epoch 24 --nboot 3 --recording 100 --dtype bf16 --train_all_layers \
    --skip_param_record --seed 5 --ngpu 8 ) > $O/${TAG}_nb_vae_ngpu8.log 2>&1
tr '\r' '\n' < $O/${TAG}_nb_vae_ngpu8.log | grep -a "Done training\|after the first\|real" | tail -4 | tee -a $O/${TAG}_summary.txt
for d in /sys/bus/pci/devices/*; do if [ -f $d/numa_node ] && grep -qi "0x10de" $d/vendor 2>/dev/null; then echo "$d $(cat $d/numa_node) $(cat $d/class)"; fi; done | head -12 >> $O/${TAG}_summary.txt
