"""Data-parallel correctness on >= 2 GPUs (skipped on a single-GPU box): G-GPU step == 1-GPU step on the
concatenated batch, replicas bit-identical."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("bf16", ["0", "1"])
def test_data_parallel_matches_single_gpu(bf16):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, DP_BF16=bf16)
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541",
                        os.path.join(ROOT, "tests", "dp_worker.py")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "dp ok" in p.stdout
