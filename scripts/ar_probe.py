"""Stand-alone NCCL all-reduce bandwidth at the gradient bucket sizes (torchrun, one rank per GPU)."""
import os, sys, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
for mb in (122.9, 38.8, 8.0, 1.0):
    n = int(mb * 1e6 / 4)
    x = torch.ones(n, device="cuda")
    for _ in range(5): dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(20): dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    if rank == 0: print("ar_probe %s: %.1f MB  %.3f ms  algbw %.0f GB/s" % (os.environ.get("TAG", ""), mb, ms, mb / ms), flush=True)
dist.destroy_process_group()
