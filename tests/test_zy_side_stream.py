"""The "side_stream" option (model.cu: side_fork / side_join; DESIGN.md section 4, "Side stream") only changes WHERE
optimiser-only work of a backward pass and the first layer's effective bias are issued, never what is computed.

Tolerances, measured on a B200 with scripts/side_diag.py (same step run off / off / on / on):
* fp32 mode: every gradient agrees to float-atomic order noise (<= 3e-7 observed, off vs off the same);
* bf16 mode: the effective bias is a sum of float atomics, and next to the panel build its order varies from run to run
  (one stream: repeatable in practice).  The 1e-7 that moves in the logits flips the bf16 rounding of a few hundred of
  the B x D gradient entries K4 / K5 write, so every tensor DOWNSTREAM of that bf16 array moves by 1e-5 ... 1e-4 of its
  norm — between two runs with the option on as much as between on and off — while the tensors K4 feeds in fp32 (the nu
  path, depth) stay at 1e-7.  Far inside the mode's 1e-2; a missing join would show as O(1)."""
import importlib

import numpy as np
import pytest

from helpers import rel_err
from test_gpu_edge import make_model, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mmv():
    return importlib.import_module("mm-vae_b200")


@pytest.fixture(scope="module")
def ctx(mmv):
    c = mmv.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("dtype", ["bf16", "f32"])
@pytest.mark.parametrize("model", ["nb1", "vmf"])
def test_side_stream_on_and_off_agree(mmv, ctx, model, dtype):
    D, N, B = 8000, 400, 256
    tol = 1e-3 if dtype == "bf16" else 1e-5
    rng = np.random.default_rng(77)
    x = synth(rng, N, D, empty_every=13)
    batches = [rng.choice(N, size=B, replace=False) for _ in range(3)]
    res = []
    for side in (0, 1):
        m, neps = make_model(mmv, ctx, model, D, [256, 128], [128, 256], 16, dtype, np.random.default_rng(3))
        m.set_option("side_stream", side)
        csr = mmv.CSR.from_dense(ctx, x)
        out = []
        for i, cells in enumerate(batches):      # consecutive steps on one model: the join at a step's end orders the next
            eps = [np.random.default_rng(100 + 10 * i + j).normal(size=(B, n)).astype(np.float32) for j, n in enumerate(neps)]
            t = m.step(csr, cells, eps=eps, kl_weight=0.7, flags=mmv.FWD | mmv.BWD)
            out.append((t["loss"], {k: m.get_grad(k).copy() for k in m.param_info()}))
        t = m.step(csr, batches[0], eps=eps, kl_weight=0.7, flags=mmv.FWD | mmv.BWD | mmv.OPTIM)
        assert np.isfinite(t["loss"])
        res.append(out)
        m.close()
        csr.close()
    for (l0, g0), (l1, g1) in zip(*res):
        assert abs(l0 - l1) <= 1e-5 * abs(l0)
        worst = {k: rel_err(g1[k], g0[k]) for k in g0}
        assert max(worst.values()) < tol, {k: v for k, v in worst.items() if v >= tol}
