"""Data-parallel correctness on >= 2 GPUs (skipped on a single-GPU box): G-GPU step == 1-GPU step on the
concatenated batch, replicas bit-identical."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", ["small", "wide"])
@pytest.mark.parametrize("bf16", ["0", "1"])
def test_data_parallel_matches_single_gpu(bf16, case):
    """case "wide" = SURVEY cfg-4 widths (D = 30 000, [1024, 256] / [256, 1024], 256 cells per rank): the bucketed
    first-layer backward, the persistent decoder-last dW GEMM, the grouped leftovers and the un-fused clip norm are
    all on the tested path; parameters AND gradients against the 1-GPU step on the concatenated batch, every replica
    tensor bit-identical, per-rank noise streams distinct."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, DP_BF16=bf16, DP_CASE=case)
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541",
                        os.path.join(ROOT, "tests", "dp_worker.py")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "dp ok" in p.stdout
