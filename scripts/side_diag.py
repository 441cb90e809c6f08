"""One-off diagnostic: gradients of the same step with the side stream off / off again / on (run-to-run noise vs the option)."""
import importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from helpers import rel_err
import test_gpu_edge as E
mmv = importlib.import_module("mm-vae_b200")
ctx = mmv.Context(0)
D, N, B = 8000, 400, 256
rng = np.random.default_rng(77)
x = E.synth(rng, N, D, empty_every=13)
cells = rng.choice(N, size=B, replace=False)
for model, dtype in (("nb1", "bf16"), ("vmf", "bf16"), ("nb1", "f32")):
    res = []
    for side in (0, 0, 1, 1):
        m, neps = E.make_model(mmv, ctx, model, D, [256, 128], [128, 256], 16, dtype, np.random.default_rng(3))
        m.set_option("side_stream", side)
        csr = mmv.CSR.from_dense(ctx, x)
        eps = [np.random.default_rng(100 + j).normal(size=(B, n)).astype(np.float32) for j, n in enumerate(neps)]
        t = m.step(csr, cells, eps=eps, kl_weight=0.7, flags=mmv.FWD | mmv.BWD)
        res.append((t["loss"], {k: m.get_grad(k).copy() for k in m.param_info()}))
        m.close(); csr.close()
    print(model, dtype, "loss", [r[0] for r in res])
    for k in res[0][1]:
        a = [rel_err(res[i][1][k], res[0][1][k]) for i in (1, 2, 3)]
        if max(a) > 1e-7:
            print(f"  {k:28s} off/off {a[0]:.2e}  on/off {a[1]:.2e}  on2/off {a[2]:.2e}  on/on {rel_err(res[3][1][k], res[2][1][k]):.2e}  |g| {np.linalg.norm(res[0][1][k]):.3e}")
