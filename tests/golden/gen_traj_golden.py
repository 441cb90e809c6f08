#!/usr/bin/env python
"""Golden TRAJECTORIES for the C++ mains (tests/test_gpu_host_main.py::test_main_trajectory_matches_reference_loop)
and the oracle (tests/test_oracle_golden.py::test_oracle_trajectory).

oracle/_ref/ref_model_harness train runs the reference's epoch x batch x bootstrap schedule
(mmvae_alg.hh:254-334) around the reference's own model / loss / Adam / clip calls with a seeded
mt19937 bootstrap stream and the eps of every forward saved in call order.  Runs in the build
container only (needs /root/reference); the .npz files are committed.

    python tests/golden/gen_traj_golden.py
"""
import glob
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_model_harness")

CASES = {
    "traj_nb": dict(model="nb", N=70, D=120, B=32, enc="24", dec="24", R=4, H=1, Rnu=1, epochs=3, nboot=2,
                    seed=5, train_all=1, kl_min=1e-2, kl_max=1.0, kl_discount=0.1, lr=1e-3, density=0.15),
    "traj_vmf": dict(model="vmf", N=60, D=96, B=25, enc="20", dec="20", R=4, epochs=3, nboot=3,
                     seed=9, train_all=0, kl_min=1e-2, kl_max=1.0, kl_discount=0.1, lr=1e-3, density=0.2),
}


def main():
    if not os.path.exists(HARNESS):
        sys.exit("build the harness first: make -C oracle")
    for name, cfg in CASES.items():
        rng = np.random.default_rng(cfg["seed"])
        x = ((rng.random((cfg["N"], cfg["D"])) < cfg["density"]) * (1 + rng.poisson(1.5, (cfg["N"], cfg["D"])))).astype(np.float32)
        for i in range(cfg["N"]):
            if x[i].sum() == 0:
                x[i, rng.integers(0, cfg["D"])] = 1
        x[3, 7] = 41
        tmp = tempfile.mkdtemp(prefix="traj_")
        try:
            x.tofile(os.path.join(tmp, "x.bin"))
            cmd = [HARNESS, "train", "--threads", "1", "--out", os.path.join(tmp, "out"), "--x", os.path.join(tmp, "x.bin")]
            for k, v in cfg.items():
                if k != "density":
                    cmd += ["--" + k, str(v)]
            subprocess.run(cmd, check=True, stderr=subprocess.DEVNULL)
            arrays = {"x": x, "config": np.frombuffer(json.dumps(cfg).encode(), dtype=np.uint8)}
            for f in sorted(glob.glob(os.path.join(tmp, "out", "*.npy"))):
                arrays[os.path.basename(f)[:-4]] = np.load(f)
            np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrays)
            print(name, "scores", arrays["scores"], f"{os.path.getsize(os.path.join(HERE, name + '.npz')) / 1024:.0f} KiB")
        finally:
            shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
