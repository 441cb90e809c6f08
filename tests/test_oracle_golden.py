"""The oracle (oracle/*.py) against the golden vectors produced by the
reference itself (tests/golden/, made by oracle/_ref harnesses).  CPU only."""
import gzip
import os

import numpy as np
import pytest
import torch

from oracle import bgzf_mtx, model_oracle as mo
from helpers import GOLDEN, MODEL_CASES, Golden, rel_err, to_t

torch.set_num_threads(1)


@pytest.mark.parametrize("case", MODEL_CASES)
def test_model_oracle_matches_reference(case):
    """Forward outputs, loss terms, grads, grad-norm and the Adam trajectory."""
    g = Golden(case)
    cfg = g.cfg
    model = cfg["model"]
    P = to_t(g.params("init"))
    kw = dict(kap_min=cfg.get("kappa_min", 0.1), kap_max=cfg.get("kappa_max", 10.0), relu=bool(cfg.get("relu", 0)))
    adam = mo.Adam(lr=1e-3, weight_decay=1e-4)
    train_all = bool(cfg.get("train_all", 0))
    acc = {}  # grads of the unregistered Sequentials accumulate (zero_grad misses them)
    for s in range(cfg["steps"]):
        x, c = g.batch(s)
        eps = [torch.from_numpy(e) for e in g.eps(s)]
        terms, outs, grads = mo.run_step(model, P, torch.from_numpy(x), torch.from_numpy(c), eps,
                                         cfg["kl"], **kw)
        for k, ref in g.group(f"s{s}__out").items():
            got = terms[k] if k in terms else outs[k]
            assert rel_err(got.numpy().reshape(ref.shape), ref) < 1e-6, (case, s, k)
        registered = set(g.group(f"s{s}__grad").keys())
        frozen = _unregistered(model, P) if not train_all else set()
        gref = g.group(f"s{s}__grad")
        opt_grads = {}
        for k, gr in grads.items():
            if k in frozen:
                acc[k] = acc.get(k, 0) + gr
                assert rel_err(acc[k].numpy(), gref[k]) < 1e-5, (case, s, k)
            else:
                assert rel_err(gr.numpy(), gref[k]) < 1e-5, (case, s, k)
                opt_grads[k] = gr.clone()
        assert registered == set(grads.keys())
        gn = mo.clip_grad_norm_(list(opt_grads.values()), 1.0)
        assert abs(gn - float(g.a[f"s{s}__gradnorm"][0])) <= 1e-5 * max(1.0, gn)
        adam.step(P, opt_grads)
        for k, ref in g.params(f"s{s}").items():
            assert rel_err(P[k].numpy(), ref) < 1e-6, (case, s, k)


def _unregistered(model, P):
    """Reference quirk (nb.hh:335-379, vmf.hh:341-385): encoder/decoder
    Sequentials are never register_module'd -> not optimised, grads accumulate."""
    if model == "nb":
        pre = ("mu_encoding", "mu_decoding")
    else:
        pre = ("encoding", "decoding", "0.")
    return {k for k in P if k.startswith(pre)}


@pytest.mark.parametrize("case", MODEL_CASES)
def test_encode_recorder_path(case):
    """Eval-mode encode without covariates (nb.hh:419-431, vmf.hh:267-281)."""
    g = Golden(case)
    P = to_t(g.params("init"))
    x = torch.from_numpy(g.a["x"])
    mean, lnvar = (mo.nb_encode_mu if g.cfg["model"] == "nb" else mo.vmf_encode)(P, x, relu=bool(g.cfg.get("relu", 0)))
    assert rel_err(mean.numpy(), g.a["enc__mean"]) < 1e-6
    assert rel_err(lnvar.numpy(), g.a["enc__lnvar"]) < 1e-6


def test_fastapprox_constants():
    """fasterlog(2pi), fasterlgamma(df+1)."""
    assert abs(mo.fasterlog(2 * np.pi) - 1.82167) < 1e-4
    assert abs(mo.fasterlgamma(1000.0) - 5936.51) < 0.05


# ------------------------------------------------------------------ reader
@pytest.fixture(scope="module")
def rg():
    z = np.load(os.path.join(GOLDEN, "reader_golden.npz"))
    return {k: z[k] for k in z.files}


FILES = {"a": "reader_a.mtx.gz", "b": "reader_b.mtx.gz", "ones": "reader_ones.mtx.gz"}


def test_take_float_bit_exact(rg):
    toks = bytes(rg["float_tokens"]).split(b"\n")
    got = np.array([bgzf_mtx.take_float(t) for t in toks], dtype=np.float32)
    assert got.view(np.uint32).tolist() == rg["float_values"].view(np.uint32).tolist()


@pytest.mark.parametrize("tag", list(FILES))
def test_reader_index_and_header(rg, tag):
    path = os.path.join(GOLDEN, FILES[tag])
    assert bgzf_mtx.bgzf_is_bgzf(path)
    assert list(bgzf_mtx.peek_bgzf_header(path)) == rg[f"{tag}__header"].tolist()
    idx = bgzf_mtx.build_index(path)
    text = "".join(f"{c} {o}\n" for c, o in idx).encode()
    assert text == bytes(rg[f"{tag}__index_text"])
    tmp = os.path.join(os.environ.get("TMPDIR", "/tmp"), f"oracle_{tag}_{os.getpid()}.index")
    bgzf_mtx.write_index(tmp, idx)
    try:
        assert gzip.open(tmp, "rb").read() == text
        assert bgzf_mtx.read_index(tmp).tolist() == rg[f"{tag}__table"].tolist()
    finally:
        os.remove(tmp)


@pytest.mark.parametrize("tag", list(FILES))
@pytest.mark.parametrize("blk", ["all", "mid"])
def test_reader_block_triplets(rg, tag, blk):
    path = os.path.join(GOLDEN, FILES[tag])
    beg, end = rg[f"{tag}__block_{blk}_range"].tolist()
    tr = bgzf_mtx.visit_bgzf_block(path, beg, end)
    assert [t[0] for t in tr] == rg[f"{tag}__block_{blk}_rows"].tolist()
    assert [t[1] for t in tr] == rg[f"{tag}__block_{blk}_cols"].tolist()
    vals = np.array([t[2] for t in tr], dtype=np.float32)
    assert vals.view(np.uint32).tolist() == rg[f"{tag}__block_{blk}_vals"].view(np.uint32).tolist()


@pytest.mark.parametrize("tag", list(FILES))
def test_reader_dense_minibatches(rg, tag):
    path = os.path.join(GOLDEN, FILES[tag])
    tab = rg[f"{tag}__table"]
    D = int(rg[f"{tag}__header"][0])
    for cols, dense in zip(rg[f"{tag}__batches"], rg[f"{tag}__dense"]):
        got = bgzf_mtx.read_batch_dense(path, tab, cols.tolist(), D)
        assert np.array_equal(got.view(np.uint32), dense.view(np.uint32))


@pytest.mark.parametrize("case", ["traj_nb", "traj_vmf"])
def test_oracle_trajectory(case):
    """The epoch x batch x bootstrap schedule (mmvae_alg.hh:254-334) replayed by the oracle on the reference's own
    eps / bootstrap streams (tests/golden/gen_traj_golden.py): per-batch losses, epoch scores, final parameters."""
    g = Golden(case)
    cfg = g.cfg
    model, N, B, R = cfg["model"], cfg["N"], cfg["B"], cfg["R"]
    n_eps = [R, cfg["Rnu"]] if model == "nb" else [R]
    P = to_t(g.params("init"))
    x_all = torch.from_numpy(g.a["x"])
    eps_stream, ridx = g.a["eps"], g.a["ridx"]
    adam = mo.Adam(lr=cfg["lr"], weight_decay=1e-4)
    frozen = _unregistered(model, P) if not cfg["train_all"] else set()
    pos, kb, batch_loss, scores = 0, 0, [], []

    def next_eps():
        nonlocal pos
        out = []
        for n in n_eps:
            out.append(torch.from_numpy(eps_stream[pos:pos + B * n].reshape(B, n)))
            pos += B * n
        return out

    nbatch = -(-N // B)
    for epoch in range(cfg["epochs"]):
        w = max(cfg["kl_max"] * float(np.exp(np.float32(-cfg["kl_discount"] * epoch))), cfg["kl_min"])
        tot = 0.0
        for b in range(nbatch):
            cells = (b * B + np.arange(B)) % N
            x, c = x_all[cells], torch.ones(B, 1)
            terms, _, _ = mo.run_step(model, P, x, c, next_eps(), w)
            batch_loss.append(float(terms["loss"]))
            tot += float(terms["loss"]) * B
            for _ in range(cfg["nboot"]):
                r = torch.from_numpy(ridx[kb]); kb += 1
                _, _, grads = mo.run_step(model, P, x[r], c[r], next_eps(), w)
                opt = {k: v.clone() for k, v in grads.items() if k not in frozen}
                mo.clip_grad_norm_(list(opt.values()), 1.0)
                adam.step(P, opt)
        scores.append(tot / (B * nbatch))
    assert pos == len(eps_stream) and kb == len(ridx)
    assert np.allclose(batch_loss, g.a["batch_loss"], rtol=2e-5), np.abs(np.array(batch_loss) / g.a["batch_loss"] - 1).max()
    assert np.allclose(scores, g.a["scores"], rtol=2e-5)
    for k, ref in g.params("final").items():
        assert rel_err(P[k].numpy(), ref) < 1e-4, (k, rel_err(P[k].numpy(), ref))
