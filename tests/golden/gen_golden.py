#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ from the REFERENCE itself.

Runs in the build container only (needs /root/reference and the harnesses
built by `make -C oracle`): oracle/_ref/ref_model_harness compiles the
reference's nb.hh / vmf.hh / operators.hh / angular.hh verbatim against
libtorch CPU; oracle/_ref/ref_reader_harness compiles its bgzf reader.  The
resulting small .npz / .gz files are committed; nothing on the GPU box reads
/root/reference.

    python tests/golden/gen_golden.py            # regenerate everything
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
READER = os.path.join(ROOT, "oracle", "_ref", "ref_reader_harness")

MODEL_CASES = {
    # name: config (reference ctor args nb_vae_main.cc:103-112 / vmf_vae_main.cc:100-107)
    "nb_small": dict(model="nb", D=96, B0=12, B=12, C=1, enc="24", dec="24", R=4, H=1, Rnu=1,
                     steps=3, seed=7, kl=1.0, density=0.15),
    "nb_deep": dict(model="nb", D=80, B0=16, B=10, C=3, enc="32,16", dec="16,32", R=5, H=3, Rnu=2,
                    steps=2, seed=11, kl=0.37, density=0.2),
    "nb_nohidden": dict(model="nb", D=64, B0=9, B=9, C=1, enc="-", dec="-", R=3, H=1, Rnu=1,
                        steps=2, seed=13, kl=1.0, density=0.25),
    "nb_trainall": dict(model="nb", D=72, B0=14, B=14, C=2, enc="20", dec="28", R=6, H=2, Rnu=1,
                        steps=3, seed=17, kl=0.5, density=0.2, train_all=1),
    "nb_wide": dict(model="nb", D=520, B0=6, B=6, C=1, enc="40", dec="36", R=8, H=1, Rnu=1,
                    steps=1, seed=19, kl=1.0, density=0.05),
    "vmf_small": dict(model="vmf", D=96, B0=12, B=12, C=1, enc="24", dec="24", R=4,
                      steps=3, seed=23, kl=1.0, density=0.15),
    "vmf_deep": dict(model="vmf", D=80, B0=16, B=10, C=3, enc="32,16", dec="16,32", R=5,
                     steps=2, seed=29, kl=0.37, density=0.2, ln_kappa=0.3),
    "vmf_nohidden": dict(model="vmf", D=64, B0=9, B=9, C=1, enc="-", dec="-", R=3,
                         steps=2, seed=31, kl=1.0, density=0.25, ln_kappa=1.1),
    "vmf_trainall": dict(model="vmf", D=72, B0=14, B=14, C=2, enc="20", dec="28", R=6,
                         steps=3, seed=37, kl=0.5, density=0.2, ln_kappa=0.0, train_all=1),
    "vmf_bigkappa": dict(model="vmf", D=6, B0=8, B=8, C=1, enc="5", dec="5", R=2,
                         steps=2, seed=41, kl=1.0, density=0.6, ln_kappa=1.5,
                         kappa_min=0.1, kappa_max=100.0),  # df=2 < kappa: second lbessel branch
    # --relu (nb.hh:339-345 throws with hidden layers, SURVEY 0.5: nb only without them; vmf.hh:342-343 works)
    "nb_relu": dict(model="nb", D=64, B0=11, B=11, C=1, enc="-", dec="-", R=3, H=1, Rnu=1,
                    steps=2, seed=43, kl=1.0, density=0.25, relu=1),
    "vmf_relu": dict(model="vmf", D=88, B0=14, B=12, C=2, enc="28,12", dec="12,28", R=5,
                     steps=3, seed=47, kl=0.6, density=0.2, ln_kappa=0.2, relu=1),
    "vmf_relu_trainall": dict(model="vmf", D=72, B0=13, B=13, C=1, enc="24", dec="20", R=4,
                              steps=2, seed=53, kl=1.0, density=0.2, ln_kappa=0.5, relu=1, train_all=1),
}


def synth_counts(rng, B, D, density):
    """Small non-negative integer counts, every cell non-empty; a few large values."""
    x = (rng.random((B, D)) < density) * (1 + rng.poisson(1.5, (B, D)))
    x[rng.integers(0, B), rng.integers(0, D)] = 57
    for i in range(B):
        if x[i].sum() == 0:
            x[i, rng.integers(0, D)] = 1
    return x.astype(np.float32)


def gen_model_case(name, cfg):
    rng = np.random.default_rng(cfg["seed"])
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        x = synth_counts(rng, cfg["B0"], cfg["D"], cfg["density"])
        if cfg["C"] == 1:
            c = np.ones((cfg["B0"], 1), np.float32)  # create_ones_like default (nb_vae_main.cc:68-73)
        else:
            c = rng.normal(size=(cfg["B0"], cfg["C"])).astype(np.float32)
        ridx = rng.integers(0, cfg["B0"], size=(cfg["steps"], cfg["B"])).astype(np.int64)
        x.tofile(os.path.join(tmp, "x.bin"))
        c.tofile(os.path.join(tmp, "c.bin"))
        ridx.tofile(os.path.join(tmp, "ridx.bin"))
        cmd = [HARNESS, "dump", "--threads", "1", "--out", os.path.join(tmp, "out"),
               "--x", os.path.join(tmp, "x.bin"), "--c", os.path.join(tmp, "c.bin"),
               "--ridx", os.path.join(tmp, "ridx.bin")]
        for k, v in cfg.items():
            if k != "density":
                cmd += ["--" + k, str(v)]
        subprocess.run(cmd, check=True, stderr=subprocess.DEVNULL)
        arrays = {"x": x, "c": c, "ridx": ridx,
                  "config": np.frombuffer(json.dumps(cfg).encode(), dtype=np.uint8)}
        for f in sorted(glob.glob(os.path.join(tmp, "out", "*.npy"))):
            arrays[os.path.basename(f)[:-4]] = np.load(f)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrays)
        print(f"{name}: {len(arrays)} arrays, "
              f"{os.path.getsize(os.path.join(HERE, name + '.npz')) / 1024:.0f} KiB")
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main():
    if not os.path.exists(HARNESS):
        sys.exit("build the harness first: make -C oracle")
    only = [a for a in sys.argv[1:] if not a.startswith("-")]   # optional: regenerate the named cases only
    for name, cfg in MODEL_CASES.items():
        if only and name not in only:
            continue
        gen_model_case(name, cfg)
    if only:
        return
    if os.path.exists(READER):
        sys.path.insert(0, HERE)
        import gen_reader_golden
        gen_reader_golden.main()


if __name__ == "__main__":
    main()
