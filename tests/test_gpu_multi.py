"""Data-parallel correctness on >= 2 GPUs (skipped on a single-GPU box): G-GPU step == 1-GPU step on the
concatenated batch, replicas bit-identical."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("bf16,case,wire", [("0", "small", "1"), ("1", "small", "1"), ("0", "wide", "1"), ("1", "wide", "1"),
                                            ("1", "wide", "0"), ("2", "wide", "1")])
def test_data_parallel_matches_single_gpu(bf16, case, wire):
    """case "wide" = SURVEY cfg-4 widths (D = 30 000, [1024, 256] / [256, 1024], 256 cells per rank): the bucketed
    first-layer backward, the persistent decoder-last dW GEMM, the grouped leftovers and the un-fused clip norm are
    all on the tested path; parameters AND gradients against the 1-GPU step on the concatenated batch, every replica
    tensor bit-identical, per-rank noise streams distinct.  bf16 = the "dtype" option: "2" is the fp32 tolerance on the
    tensor cores (operands as bf16 pairs; gradients reduced in fp32), held to the fp32 budget."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    # wire = "1": in bf16 mode the decoder-last and first-layer weight gradients cross NVLink as bf16 (comm_bf16, the
    # default); "0": every gradient reduced in fp32
    env = dict(os.environ, DP_BF16=bf16, DP_CASE=case, DP_COMM_BF16=wire)
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541",
                        os.path.join(ROOT, "tests", "dp_worker.py")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "dp ok" in p.stdout
