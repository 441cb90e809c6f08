"""Data-parallel worker (launched by torchrun from test_gpu_multi.py, or with --cpu from test_dist_cpu.py).

GPU mode: rank r steps on ITS shard of a global batch with injected noise; with the NCCL allreduce inside
mmv_step the post-Adam parameters of every rank must equal a single-GPU step on the concatenated batch
(SURVEY 8e pin: sum-order tolerance), and the replicas must be bit-identical to each other.
CPU mode (--cpu, gloo): only the host-side plumbing bench.py uses for N > 1 — cell sharding, unique-id
broadcast, max-over-ranks timing — no compute.
"""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def shard_range(n_cells, world, rank):
    """Contiguous cell range of a rank: [rank * n / world, (rank + 1) * n / world)."""
    return rank * n_cells // world, (rank + 1) * n_cells // world


def main_cpu():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [bytes(range(128)) if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    assert box[0] == bytes(range(128))
    lo, hi = shard_range(1001, world, rank)
    cover = [None] * world
    dist.all_gather_object(cover, (lo, hi))
    assert cover[0][0] == 0 and cover[-1][1] == 1001 and all(cover[i][1] == cover[i + 1][0] for i in range(world - 1))
    t = torch.tensor([1.0 + rank, 10.0 - rank])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert t.tolist() == [float(world), 10.0]
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank} ok")


def main_gpu():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mmv = importlib.import_module("mm-vae_b200")
    mode = int(os.environ.get("DP_BF16", "0"))    # "dtype" option: 0 fp32 FFMA, 1 bf16 tensor cores, 2 fp32 tolerance on tensor cores
    bf16 = mode == 1
    wide = os.environ.get("DP_CASE", "small") == "wide"
    if wide:
        # SURVEY cfg-4 widths: the first-layer backward runs >= 3 waves (bucketed dW1 parts, "small last part goes with
        # the leftovers"), the decoder-last dW takes the persistent GEMM, the clip norm is un-fused (N > 1) — the code
        # path bench.py --gpus N measures
        D, Bl, R, enc, dec = 30000, 256, 32, [1024, 256], [256, 1024]
    else:
        D, Bl, R, enc, dec = 512, 96, 8, [64], [64]
    B = Bl * world
    rng = np.random.default_rng(5)                       # same global data on every rank
    x = (rng.random((B, D)) < (0.05 if wide else 0.08)) * rng.integers(1, 6, (B, D))
    x = x.astype(np.float32)
    x[:, 0] = 1.0
    eps_mu, eps_nu = rng.standard_normal((B, R)).astype(np.float32), rng.standard_normal((B, 1)).astype(np.float32)

    def make(ctx):
        m = mmv.Model.nb(ctx, D, 1, enc, dec, R, 1, 1, seed=3)
        m.set_option("train_all_layers", 1)
        m.set_option("dtype", mode)
        m.set_option("comm_bf16", int(os.environ.get("DP_COMM_BF16", "1")))   # bf16 mode: the two D x h gradients as bf16 on the wire
        if wide:   # off the symmetric initial values, and a clip that acts
            m.set_param("x_mean", np.random.default_rng(8).normal(scale=0.05, size=D).astype(np.float32))
            m.set_option("grad_clip", 0.5)
        return m

    flags = mmv.FWD | mmv.BWD | mmv.OPTIM
    nsteps = 2 if wide else 3
    # ---- distributed: this rank's rows
    ctx = mmv.Context(local)
    box = [mmv.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    ctx.comm_init(world, rank, box[0])
    lo, hi = shard_range(B, world, rank)
    csr = mmv.CSR.from_dense(ctx, x[lo:hi], None)
    m = make(ctx)
    for _ in range(nsteps):
        eps = np.concatenate([eps_mu[lo:hi].ravel(), eps_nu[lo:hi].ravel()])
        terms = m.step(csr, np.arange(hi - lo), eps=eps, kl_weight=0.7, flags=flags)
    names = list(m.param_info())
    dist_params = {k: m.get_param(k) for k in names}
    dist_grads = {k: m.get_grad(k) for k in names}
    # replicas bit-identical: EVERY tensor, parameters and (all-reduced, clipped) gradients
    for k in names:
        for what, d in (("param", dist_params), ("grad", dist_grads)):
            t = torch.from_numpy(d[k].copy()).cuda()
            ref = t.clone()
            dist.broadcast(ref, src=0)
            assert torch.equal(t, ref), f"replicas diverged on {what} {k}"
    # the library's own noise (MMV_EPS_PHILOX) must differ between ranks (same model seed, lock-step draw counter)
    m.step(csr, np.arange(hi - lo), kl_weight=0.7, flags=mmv.FWD | mmv.EPS_PHILOX)
    e = torch.from_numpy(m.output("eps_mu", (hi - lo) * R)).cuda()
    e0 = e.clone()
    dist.broadcast(e0, src=0)
    if rank != 0:
        assert not torch.equal(e, e0), "every rank drew the same reparameterisation noise"
        n_e = e.numel()   # independent N(0,1) streams: sample correlation ~ N(0, 1/n), sample std ~ N(1, 1/(2n)); 4 sigma
        assert abs(float((e * e0).mean())) < 4.0 / n_e ** 0.5 and abs(float(e.std()) - 1.0) < 4.0 / (2 * n_e) ** 0.5
    # ---- single GPU on the concatenated batch (own context without a communicator)
    ctx1 = mmv.Context(local)
    csr1 = mmv.CSR.from_dense(ctx1, x, None)
    m1 = make(ctx1)
    for _ in range(nsteps):
        eps = np.concatenate([eps_mu.ravel(), eps_nu.ravel()])
        terms1 = m1.step(csr1, np.arange(B), eps=eps, kl_weight=0.7, flags=flags)
    tol = 2e-2 if bf16 else 2e-5
    assert abs(terms["loss"] - terms1["loss"]) <= tol * abs(terms1["loss"]), (terms, terms1)
    assert abs(terms["grad_norm"] - terms1["grad_norm"]) <= tol * abs(terms1["grad_norm"]), (terms, terms1)
    worst, worst_g, wk = 0.0, 0.0, ""
    for k in names:
        a, b = dist_params[k], m1.get_param(k)
        den = np.linalg.norm(b) + 1e-12
        worst = max(worst, float(np.linalg.norm(a - b) / den))
        ga, gb = dist_grads[k], m1.get_grad(k)
        if np.linalg.norm(gb) > 1e-12:
            e = float(np.linalg.norm(ga - gb) / np.linalg.norm(gb))
            if e > worst_g:
                worst_g, wk = e, k
    assert worst <= tol, f"data-parallel parameters differ from the single-GPU step: {worst}"
    # gradients of the LAST step: the all-reduced mean over ranks against the single-GPU gradient of the whole batch
    gtol = 2e-2 if bf16 else 1e-4
    assert worst_g <= gtol, f"data-parallel gradient {wk} differs from the single-GPU step: {worst_g}"
    dist.barrier()
    if rank == 0:
        print(f"dp ok: world {world}, case {'wide' if wide else 'small'}, dtype {mode}, loss {terms['loss']:.6f} vs {terms1['loss']:.6f}, "
              f"worst rel param diff {worst:.2e}, worst rel grad diff {worst_g:.2e} ({wk})")
    m.close(); m1.close(); csr.close(); csr1.close(); ctx.close(); ctx1.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main_cpu() if "--cpu" in sys.argv else main_gpu()
