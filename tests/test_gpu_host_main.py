"""The C++ mains (mm-vae_b200/host/bin/nb_vae, vmf_vae) end to end on a B200: reference command line,
bgzf-indexed Matrix Market input, resident CSR, the epoch / batch / bootstrap schedule of
mmvae_alg.hh:254-334 through the C ABI, ${out}.scores.gz and the latent tables out."""
import gzip
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
HOST = os.path.join(ROOT, "mm-vae_b200", "host")

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bins():
    subprocess.run(["make", "-C", os.path.join(ROOT, "mm-vae_b200")], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["make", "-C", HOST], check=True, stdout=subprocess.DEVNULL)
    return os.path.join(HOST, "bin")


def scores(path):
    return np.array([float(v) for v in gzip.open(path, "rt").read().split()], dtype=np.float64)


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_nb_vae_main_trains(bins, tmp_path, dtype):
    out = tmp_path / "nb"
    idx = tmp_path / "b.index"
    cmd = [os.path.join(bins, "nb_vae"), "--mtx", os.path.join(GOLDEN, "reader_b.mtx.gz"), "--idx", str(idx), "--out", str(out),
           "--batch_size", "100", "--mean_encoding", "64", "--mean_decoding", "64", "--mean_latent", "8", "--max_epoch", "6",
           "--recording", "3", "--nboot", "2", "--seed", "5", "--dtype", dtype, "--train_all_layers", "--lr", "0.01"]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    assert idx.exists()                                  # built on demand like nb_vae_main.cc:56-57
    s = scores(str(out) + ".scores.gz")
    assert s.shape == (6,) and np.all(np.isfinite(s)) and s[-1] < s[0]
    for tag in ("2", "5"):   # update_on_epoch(model, epoch) with the 0-based epoch, zero-padded to len("6") (nb.hh:590)
        m = np.loadtxt(gzip.open(f"{out}_{tag}.mu_mean.gz", "rt"))
        lv = np.loadtxt(gzip.open(f"{out}_{tag}.mu_lnvar.gz", "rt"))
        assert m.shape == (500, 8) and lv.shape == (500, 8) and np.all(np.isfinite(m))
        assert np.all(lv >= -4.0001) and np.all(lv <= 4.0001)          # clamp of nb.hh:415
    # per-parameter dumps (nb.hh:599-615 through write_tensor, mmvae_io.hh:11-28): 2-D one row per line, biases one value per line
    D = 300
    w1 = np.loadtxt(gzip.open(f"{out}_5_mu_encoding_1.weight.gz", "rt"))
    assert w1.shape == (64, D) and np.all(np.isfinite(w1))
    b1 = np.loadtxt(gzip.open(f"{out}_5_mu_encoding_1.bias.gz", "rt"))
    assert b1.shape == (64,)
    xm = np.loadtxt(gzip.open(f"{out}_5_x_mean.gz", "rt"), ndmin=2)
    assert xm.shape == (1, D)
    wd = np.loadtxt(gzip.open(f"{out}_5_mu_decoding.weight.gz", "rt"))
    assert wd.shape == (D, 64)
    assert not np.allclose(np.loadtxt(gzip.open(f"{out}_2_mu_decoding.weight.gz", "rt")), wd)   # training moved it


def test_vmf_vae_main_trains(bins, tmp_path):
    out = tmp_path / "vmf"
    cmd = [os.path.join(bins, "vmf_vae"), "--mtx", os.path.join(GOLDEN, "reader_b.mtx.gz"), "--idx", str(tmp_path / "b.index"),
           "--out", str(out), "--batch", "128", "--encoding", "32", "--decoding", "32", "--latent", "4", "--epoch", "4",
           "--recording", "4", "--seed", "3", "--train_all_layers", "--skip_param_record"]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    s = scores(str(out) + ".scores.gz")
    assert s.shape == (4,) and np.all(np.isfinite(s)) and s[-1] <= s[0]
    assert np.loadtxt(gzip.open(f"{out}_3.latent_mean.gz", "rt")).shape == (500, 4)
    assert not os.path.exists(f"{out}_3_z_decoding.weight.gz")          # --skip_param_record


def test_vmf_vae_main_with_vmf_latent(bins, tmp_path):
    """--vmf_latent (not in the reference): the latent table holds unit vectors, the second table -log kappa."""
    out = tmp_path / "vmfl"
    cmd = [os.path.join(bins, "vmf_vae"), "--mtx", os.path.join(GOLDEN, "reader_b.mtx.gz"), "--idx", str(tmp_path / "b.index"),
           "--out", str(out), "--batch", "128", "--encoding", "32", "--decoding", "32", "--latent", "6", "--epoch", "4",
           "--recording", "4", "--seed", "3", "--train_all_layers", "--skip_param_record", "--vmf_latent"]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    s = scores(str(out) + ".scores.gz")
    assert s.shape == (4,) and np.all(np.isfinite(s))
    mu = np.loadtxt(gzip.open(f"{out}_3.latent_mean.gz", "rt"))
    lv = np.loadtxt(gzip.open(f"{out}_3.latent_lnvar.gz", "rt"))
    assert mu.shape == (500, 6) and np.abs(np.linalg.norm(mu, axis=1) - 1).max() < 1e-4
    assert np.all(lv <= 0) and np.abs(lv - lv[:, :1]).max() == 0          # -log kappa, kappa >= 1, one value per cell


def test_nb_vae_main_with_csr_cache(bins, tmp_path):
    """--csr_cache: the second run trains from ${idx}.csr instead of the text and reports the same epoch losses."""
    import shutil
    mtx = tmp_path / "b.mtx.gz"
    shutil.copy(os.path.join(GOLDEN, "reader_b.mtx.gz"), mtx)
    res = []
    for k in range(2):
        out = tmp_path / f"run{k}"
        cmd = [os.path.join(bins, "nb_vae"), "--mtx", str(mtx), "--out", str(out), "--batch_size", "100", "--mean_encoding", "32",
               "--mean_decoding", "32", "--mean_latent", "4", "--max_epoch", "3", "--recording", "3", "--nboot", "1", "--seed", "5",
               "--skip_param_record", "--csr_cache"]
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
        assert p.returncode == 0, p.stderr[-2000:]
        assert ("binary CSR cache" in p.stderr) == (k == 1)
        res.append(scores(str(out) + ".scores.gz"))
    assert os.path.exists(str(mtx) + ".index.csr")
    assert np.allclose(res[0], res[1], rtol=1e-4)


def test_main_fails_like_reference_without_out(bins):
    p = subprocess.run([os.path.join(bins, "nb_vae"), "--mtx", os.path.join(GOLDEN, "reader_b.mtx.gz")],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60)
    assert p.returncode == 1 and "need output file header" in p.stderr


def _write_mtx(bins, x, path):
    """cells x genes dense counts -> bgzf Matrix Market (genes x cells, sorted by cell = by column) through the repo's writer."""
    cell, gene = np.nonzero(x)
    txt = str(path) + ".txt"
    with open(txt, "w") as f:
        f.write(f"%%MatrixMarket matrix coordinate integer general\n{x.shape[1]} {x.shape[0]} {len(cell)}\n")
        for c, g in zip(cell, gene):
            f.write(f"{g + 1} {c + 1} {int(x[c, g])}\n")
    subprocess.run([os.path.join(bins, "mmv_reader_tool"), "write", txt, str(path)], check=True)


@pytest.mark.parametrize("case,dtype", [("traj_nb", "f32"), ("traj_vmf", "f32"), ("traj_nb", "bf16")])
def test_main_trajectory_matches_reference_loop(bins, tmp_path, case, dtype):
    """SURVEY 4 integration row: the C++ main (option parsing, reader, train_vae_model, mmv_step, score file) against
    the reference's own loop (mmvae_alg.hh:259-321 around the reference's model / loss / Adam, driven by
    oracle/ref_model_harness train -> tests/golden/traj_*.npz) on the same data, initial parameters, bootstrap stream
    (std::mt19937(seed) + uniform_int_distribution, :242-244) and injected eps: epoch losses in ${out}.scores.gz."""
    import json
    z = np.load(os.path.join(GOLDEN, case + ".npz"))
    cfg = json.loads(bytes(z["config"]).decode())
    mtx = tmp_path / "x.mtx.gz"
    _write_mtx(bins, z["x"], mtx)
    init = tmp_path / "init"
    init.mkdir()
    for k in z.files:
        if k.startswith("init__param__"):
            z[k].astype(np.float32).tofile(init / (k[len("init__param__"):] + ".f32"))
    z["eps"].astype(np.float32).tofile(tmp_path / "eps.f32")
    out = tmp_path / "o"
    common = ["--mtx", str(mtx), "--out", str(out), "--batch_size", str(cfg["B"]), "--max_epoch", str(cfg["epochs"]),
              "--recording", "100", "--nboot", str(cfg["nboot"]), "--seed", str(cfg["seed"]), "--lr", str(cfg["lr"]),
              "--kl_min", str(cfg["kl_min"]), "--kl_max", str(cfg["kl_max"]), "--kl_discount", str(cfg["kl_discount"]),
              "--init_dir", str(init), "--eps_file", str(tmp_path / "eps.f32"), "--dtype", dtype, "--skip_param_record"]
    if cfg["train_all"]:
        common.append("--train_all_layers")
    if cfg["model"] == "nb":
        cmd = [os.path.join(bins, "nb_vae"), *common, "--mean_encoding", cfg["enc"], "--mean_decoding", cfg["dec"],
               "--mean_latent", str(cfg["R"]), "--overdisp_encoding", str(cfg["H"]), "--overdisp_latent", str(cfg["Rnu"])]
    else:
        cmd = [os.path.join(bins, "vmf_vae"), *common, "--encoding", cfg["enc"], "--decoding", cfg["dec"], "--latent", str(cfg["R"])]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    nparam = sum(k.startswith("init__param__") for k in z.files)
    assert f"Initial values of {nparam} / {nparam} parameters" in p.stderr, p.stderr[-1500:]
    s = scores(str(out) + ".scores.gz")
    tol = 3e-4 if dtype == "f32" else 1e-2
    print(case, dtype, "scores", s, "reference", z["scores"], "max rel", np.abs(s / z["scores"] - 1).max())
    assert s.shape == z["scores"].shape and np.allclose(s, z["scores"], rtol=tol), (s, z["scores"])


def test_nb_vae_main_two_gpus(bins, tmp_path):
    """--ngpu 2: one process, a host thread + context per GPU, per-shard ingest / upload, in-process NCCL communicator.
    The run must train (finite, decreasing loss) and fill the whole latent table; with the same seed the global minibatch
    differs from the 1-GPU schedule (B cells per GPU), so the check is on behaviour, not on equality with --ngpu 1."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = tmp_path / "nb2"
    cmd = [os.path.join(bins, "nb_vae"), "--mtx", os.path.join(GOLDEN, "reader_b.mtx.gz"), "--idx", str(tmp_path / "b.index"),
           "--out", str(out), "--batch_size", "64", "--mean_encoding", "64", "--mean_decoding", "64", "--mean_latent", "8",
           "--max_epoch", "6", "--recording", "3", "--nboot", "2", "--seed", "5", "--train_all_layers", "--lr", "0.01",
           "--skip_param_record", "--ngpu", "2"]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    assert "shard [0, 250)" in p.stderr and "shard [250, 500)" in p.stderr
    s = scores(str(out) + ".scores.gz")
    assert s.shape == (6,) and np.all(np.isfinite(s)) and s[-1] < s[0]
    m = np.loadtxt(gzip.open(f"{out}_5.mu_mean.gz", "rt"))
    assert m.shape == (500, 8) and np.all(np.isfinite(m)) and np.abs(m[:250]).sum() > 0 and np.abs(m[250:]).sum() > 0


@pytest.mark.parametrize("model", ["nb", "vmf"])
def test_tu_written_against_reference_names_trains_identically(bins, tmp_path, model):
    """tests/refcompat/refcompat_tu.cc — reference header names, strong dimension types, a LOSS functor returning
    torch::Tensor, the six-argument train_vae_model (mmvae_alg.hh:200-213) — built against host/include/refcompat and the real
    library: the same epoch losses and latent table as this repo's own main on the same command line (same seeds)."""
    lib = os.path.join(ROOT, "mm-vae_b200")
    exe = tmp_path / "tu_bin"
    cmd = ["g++", "-std=c++17", "-O2", "-pthread", *(["-DREFCOMPAT_VMF"] if model == "vmf" else []),
           "-I" + os.path.join(HOST, "include", "refcompat"), "-I" + os.path.join(HOST, "include"),
           os.path.join(ROOT, "tests", "refcompat", "refcompat_tu.cc"), "-o", str(exe), "-L" + lib, "-lmmvae_b200",
           "-Wl,-rpath," + lib, "-lz"]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout[-3000:]
    if model == "nb":
        args = ["--batch_size", "100", "--mean_encoding", "64", "--mean_decoding", "64", "--mean_latent", "8", "--max_epoch", "4",
                "--recording", "4", "--nboot", "2", "--seed", "5", "--train_all_layers", "--lr", "0.01", "--skip_param_record"]
        table = "3.mu_mean.gz"
    else:
        args = ["--batch", "128", "--encoding", "32", "--decoding", "32", "--latent", "4", "--epoch", "4", "--recording", "4",
                "--seed", "3", "--train_all_layers", "--skip_param_record"]
        table = "3.latent_mean.gz"
    res = {}
    for tag, prog in (("own", os.path.join(bins, "nb_vae" if model == "nb" else "vmf_vae")), ("tu", str(exe))):
        d = tmp_path / tag
        d.mkdir()
        c = [prog, "--mtx", os.path.join(GOLDEN, "reader_b.mtx.gz"), "--idx", str(d / "b.index"), "--out", str(d / "o"), *args]
        p = subprocess.run(c, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
        assert p.returncode == 0, p.stderr[-2000:]
        res[tag] = (scores(str(d / "o") + ".scores.gz"), np.loadtxt(gzip.open(f"{d / 'o'}_{table}", "rt")))
    assert np.all(np.isfinite(res["own"][0])) and res["own"][0][-1] <= res["own"][0][0]
    # same kernels on the same inputs; the only freedom is the order of fp32 atomics inside a step
    assert np.allclose(res["own"][0], res["tu"][0], rtol=1e-4)
    assert np.allclose(res["own"][1], res["tu"][1], rtol=1e-3, atol=1e-4)
