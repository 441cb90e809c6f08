"""N > 1 host plumbing on CPU: world_size 2, gloo (the launch bench.py uses, minus the GPU)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_plumbing_gloo():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533",
                        os.path.join(ROOT, "tests", "dp_worker.py"), "--cpu"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "rank 0 ok" in p.stdout and "rank 1 ok" in p.stdout
