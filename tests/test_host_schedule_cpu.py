"""Host logic without a GPU: the C++ mains (mm-vae_b200/host/bin) run against a RECORDING TEST DOUBLE of the C ABI
(tests/stub/mmv_stub.c, put in front of the real library through LD_LIBRARY_PATH) and the log of calls is checked
against the schedule of the reference's train_vae_model (mmvae_alg.hh:254-334): per epoch and minibatch one
train-mode evaluation forward, then nboot bootstrap steps; batches wrap modulo N (:264-266); the KL weight is
max(kl_max exp(-kl_discount epoch), kl_min) (nb_vae_main.cc:27-32); the epoch loss is the mean of the evaluation
losses (:281,324); recorders fire every `recording` epochs.  The stub computes nothing and is not a fallback."""
import gzip
import math
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
HOST = os.path.join(ROOT, "mm-vae_b200", "host")
FWD, BWD, OPTIM, PHILOX, DEFERRED = 1, 2, 4, 8, 16
N, D = 500, 300            # reader_b.mtx.gz


@pytest.fixture(scope="module")
def stub(tmp_path_factory):
    if not os.path.exists(os.path.join(ROOT, "mm-vae_b200", "libmmvae_b200.so")):
        pytest.skip("the mains link against the real library; build() first")
    subprocess.run(["make", "-C", HOST], check=True, stdout=subprocess.DEVNULL)
    d = tmp_path_factory.mktemp("stub")
    subprocess.run(["gcc", "-O1", "-shared", "-fPIC", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "stub", "mmv_stub.c"), "-o", str(d / "libmmvae_b200.so"), "-lm"], check=True)
    return d


def run_main(stub, tmp_path, prog, *args):
    log = tmp_path / "calls.log"
    env = dict(os.environ, LD_LIBRARY_PATH=str(stub), MMV_STUB_LOG=str(log))
    out = tmp_path / "out"
    cmd = [os.path.join(HOST, "bin", prog), "--mtx", os.path.join(GOLDEN, "reader_b.mtx.gz"), "--idx", str(tmp_path / "b.index"),
           "--out", str(out), *map(str, args)]
    p = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert p.returncode == 0, p.stderr[-2000:]
    calls = []
    for ln in open(log):
        t = ln.split()
        calls.append((t[0], dict(kv.split("=", 1) for kv in t[1:] if "=" in kv) if t[0] != "set_option" else {"key": t[1], "value": t[2]}))
    return calls, out


def scores(path):
    return np.array([float(v) for v in gzip.open(path, "rt").read().split()])


def test_nb_vae_schedule(stub, tmp_path):
    B, E, nboot, rec = 128, 5, 2, 2
    calls, out = run_main(stub, tmp_path, "nb_vae", "--batch_size", B, "--max_epoch", E, "--recording", rec, "--nboot", nboot,
                          "--seed", 5, "--skip_param_record", "--mean_latent", 3, "--kl_discount", 0.5, "--kl_min", 0.2,
                          "--lr", 0.01, "--grad_clip", 2, "--dtype", "bf16", "--train_all_layers")
    names = [c[0] for c in calls]
    assert names[:3] == ["ctx_create", "csr_upload", "nb_model"]
    up = calls[1][1]
    assert (up["cells"], up["genes"], up["covar"], up["c_dim"]) == (str(N), str(D), "0", "1")     # all-ones covariate stays implicit
    opts = {c[1]["key"]: float(c[1]["value"]) for c in calls if c[0] == "set_option"}
    assert opts["lr"] == 0.01 and opts["grad_clip"] == 2 and opts["weight_decay"] == 1e-4 and opts["dtype"] == 1
    assert opts["train_all_layers"] == 1
    nbatch = math.ceil(N / B)
    body = [c for c in calls if c[0] in ("step", "collect", "encode")]
    i = 0
    for epoch in range(E):
        klw = max(1.0 * math.exp(-0.5 * epoch), 0.2)
        for b in range(nbatch):
            ev = body[i][1]; i += 1
            assert body[i - 1][0] == "step" and int(ev["flags"]) == FWD | PHILOX | DEFERRED and ev["out"] == "1" and ev["boot"] == "0"
            assert (int(ev["B"]), int(ev["first"]), int(ev["last"])) == (B, (b * B) % N, (b * B + B - 1) % N)      # wraps modulo N
            assert abs(float(ev["kl"]) - klw) < 1e-6 * klw and ev["eps"] == "0"
            for _ in range(nboot):
                st = body[i][1]; i += 1
                assert int(st["flags"]) == FWD | BWD | OPTIM | PHILOX and st["out"] == "0" and st["boot"] == "1"
                assert 0 <= int(st["bmin"]) <= int(st["bmax"]) < B and st["first"] == ev["first"]
                assert abs(float(st["kl"]) - klw) < 1e-6 * klw
            if (epoch + 1) % rec == 0:                       # recorder: update_on_batch encodes the minibatch
                en = body[i][1]; i += 1
                assert body[i - 1][0] == "encode" and (int(en["B"]), en["first"]) == (B, ev["first"])
        assert body[i] == ("collect", {"pending": "1"}); i += 1
    assert i == len(body)
    # epoch loss = mean of the evaluation losses (the stub's k-th evaluation reports 1000 - k), deferred or not
    want = [np.mean([1000 - (e * nbatch + b) for b in range(nbatch)]) for e in range(E)]
    assert np.allclose(scores(str(out) + ".scores.gz"), want)
    # latent tables of the recorded epochs: one row per data point, filled in data-point order through the batches
    for tag in ("1", "3"):
        m = np.loadtxt(gzip.open(f"{out}_{tag}.mu_mean.gz", "rt"))
        assert m.shape == (N, 3) and np.array_equal(m[:, 0], np.arange(N))
    assert not os.path.exists(f"{out}_0.mu_mean.gz") and not os.path.exists(f"{out}_4.mu_mean.gz")


def test_nb_vae_ngpu2_schedule(stub, tmp_path):
    """--ngpu 2 (one process, one host thread + context per GPU): contiguous shards [g N / 2, (g + 1) N / 2) ingested and
    uploaded per rank, one in-process communicator, the same schedule on every shard with shard-local cell ids, the number
    of minibatches of the largest shard, losses and files from rank 0, the latent table filled by both ranks."""
    B, E, nboot = 96, 2, 2
    calls, out = run_main(stub, tmp_path, "nb_vae", "--batch_size", B, "--max_epoch", E, "--recording", 1, "--nboot", nboot,
                          "--seed", 5, "--skip_param_record", "--mean_latent", 3, "--ngpu", 2)
    names = [c[0] for c in calls]
    assert names.count("comm_unique_id") == 1 and names.count("ctx_create") == 2 and names.count("nb_model") == 2
    ups = sorted((int(c[1]["dev"]), int(c[1]["cells"])) for c in calls if c[0] == "csr_upload")
    assert ups == [(0, N // 2), (1, N - N // 2)]
    inits = sorted((int(c[1]["device"]), int(c[1]["nranks"]), int(c[1]["rank"]), c[1]["id_ok"]) for c in calls if c[0] == "comm_init")
    assert inits == [(0, 2, 0, "1"), (1, 2, 1, "1")]
    nbatch = math.ceil(math.ceil(N / 2) / B)
    for dev in (0, 1):
        body = [c for c in calls if c[0] in ("step", "encode") and int(c[1]["dev"]) == dev]
        nloc = N // 2 if dev == 0 else N - N // 2
        i = 0
        for epoch in range(E):
            for b in range(nbatch):
                ev = body[i][1]; i += 1
                assert int(ev["flags"]) == FWD | PHILOX | DEFERRED and int(ev["ncells"]) == nloc
                assert (int(ev["first"]), int(ev["last"])) == ((b * B) % nloc, (b * B + B - 1) % nloc)     # shard-local ids, wrap
                for _ in range(nboot):
                    st = body[i][1]; i += 1
                    assert int(st["flags"]) == FWD | BWD | OPTIM | PHILOX and st["boot"] == "1"
                en = body[i]; i += 1
                assert en[0] == "encode" and en[1]["first"] == ev["first"]
        assert i == len(body)
    assert scores(str(out) + ".scores.gz").shape == (E,)
    m = np.loadtxt(gzip.open(f"{out}_1.mu_mean.gz", "rt"))
    # the stub's encode returns the (shard-local) cell id: rows of shard 1 restart at 0 -> the table is global, ids local
    assert m.shape == (N, 3)
    assert np.array_equal(m[:N // 2, 0], np.arange(N // 2)) and np.array_equal(m[N // 2:, 0], np.arange(N - N // 2))


def test_init_dir_and_eps_file_hooks(stub, tmp_path):
    """--init_dir / --eps_file (reproducibility hooks): named parameter files are loaded through mmv_model_set_param before
    training, and every forward takes its noise from the file instead of the library's Philox draws."""
    B, R = 250, 3
    init = tmp_path / "init"
    init.mkdir()
    np.zeros(D, np.float32).tofile(init / "x_mean.f32")
    nfwd = 1 * 2 * (1 + 2)
    np.arange(nfwd * B * (R + 1), dtype=np.float32).tofile(tmp_path / "eps.f32")
    # the stub exposes no parameters (num_params = 0): the loader must then load nothing and not fail
    calls, _ = run_main(stub, tmp_path, "nb_vae", "--batch_size", B, "--max_epoch", 1, "--nboot", 2, "--seed", 5, "--skip_param_record",
                        "--mean_latent", R, "--init_dir", init, "--eps_file", tmp_path / "eps.f32")
    steps = [c[1] for c in calls if c[0] == "step"]
    assert len(steps) == nfwd and all(s["eps"] == "1" for s in steps)
    assert all(int(s["flags"]) & PHILOX == 0 for s in steps)


def test_vmf_vae_schedule_and_vmf_latent_option(stub, tmp_path):
    calls, out = run_main(stub, tmp_path, "vmf_vae", "--batch", 250, "--epoch", 2, "--recording", 2, "--nboot", 1, "--latent", 4,
                          "--encoding", "16,8", "--decoding", "8", "--kappa_max", 20, "--skip_param_record", "--vmf_latent")
    mdl = [c[1] for c in calls if c[0] == "vmf_model"][0]
    assert (mdl["x"], mdl["z"], mdl["n_enc"], mdl["n_dec"], mdl["kmax"]) == (str(D), "4", "2", "1", "20")
    opts = {c[1]["key"]: float(c[1]["value"]) for c in calls if c[0] == "set_option"}
    assert opts["vmf_latent"] == 1 and opts["dtype"] == 0 and opts["train_all_layers"] == 0
    steps = [c[1] for c in calls if c[0] == "step"]
    assert len(steps) == 2 * 2 * (1 + 1)
    assert [float(s["kl"]) for s in steps[:4]] == [1.0] * 4                          # epoch 0: kl_max
    assert all(abs(float(s["kl"]) - math.exp(-0.1)) < 1e-6 for s in steps[4:])      # epoch 1: default discount 0.1
    assert np.loadtxt(gzip.open(f"{out}_1.latent_mean.gz", "rt")).shape == (N, 4)
    # without the flag the option is never touched: the default path is the reference's
    calls2, _ = run_main(stub, tmp_path, "vmf_vae", "--batch", 250, "--epoch", 1, "--latent", 4, "--skip_param_record")
    assert "vmf_latent" not in {c[1]["key"] for c in calls2 if c[0] == "set_option"}


def test_covariate_file_reaches_the_abi(stub, tmp_path):
    """--covar: a C x N Matrix Market file is read through the same reader, uploaded as the dense
    covariate block, and the model is built with c_dim = C (nb_vae_main.cc:65-82)."""
    from oracle import bgzf_mtx
    rng = np.random.default_rng(3)
    C = 2
    lines = [f"%%MatrixMarket matrix coordinate real general\n{C} {N} {C * N}\n"]
    for c in range(N):
        for r in range(C):
            lines.append(f"{r + 1} {c + 1} {rng.normal():.4f}\n")
    cov = tmp_path / "cov.mtx.gz"
    bgzf_mtx.bgzf_write(str(cov), "".join(lines).encode())
    calls, _ = run_main(stub, tmp_path, "nb_vae", "--covar", cov, "--batch_size", 250, "--max_epoch", 1, "--nboot", 1,
                        "--skip_param_record")
    up = [c[1] for c in calls if c[0] == "csr_upload"][0]
    mdl = [c[1] for c in calls if c[0] == "nb_model"][0]
    assert (up["covar"], up["c_dim"], mdl["c"]) == ("1", str(C), str(C))
    assert os.path.exists(str(cov) + ".index")                      # built on demand like the data index


def test_mains_under_sanitizers(stub, tmp_path):
    """train_vae_model, the option parsers, the covariate loader and the recorders' gz writers built with
    AddressSanitizer + UBSan and run against the test double: any finding aborts the program."""
    exe = {}
    for prog in ("nb_vae", "vmf_vae"):
        exe[prog] = str(tmp_path / (prog + "_asan"))
        cmd = ["/usr/bin/g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-pthread",
               "-I" + os.path.join(HOST, "include"), os.path.join(HOST, "src", prog + "_main.cc"), "-o", exe[prog],
               "-L" + str(stub), "-lmmvae_b200", "-lz"]
        if subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode != 0:
            pytest.skip("no sanitizer runtime for the host compiler on this machine")
    env = dict(os.environ, LD_LIBRARY_PATH=str(stub), MMV_STUB_LOG=str(tmp_path / "log"), ASAN_OPTIONS="detect_leaks=0",
               UBSAN_OPTIONS="halt_on_error=1")
    mtx = os.path.join(GOLDEN, "reader_b.mtx.gz")
    for prog, extra in (("nb_vae", ["--mean_encoding", "16,8", "--mean_decoding", "8", "--mean_latent", "3", "--csr_cache"]),
                        ("nb_vae", ["--mean_latent", "3", "--csr_cache"]),
                        ("vmf_vae", ["--latent", "4", "--vmf_latent"])):
        p = subprocess.run([exe[prog], "--mtx", mtx, "--idx", str(tmp_path / "b.index"), "--out", str(tmp_path / "o"), "--batch_size",
                            "96", "--max_epoch", "3", "--recording", "1", "--nboot", "2", *extra],
                           env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
        assert p.returncode == 0, p.stderr[-3000:]
    assert "binary CSR cache" in p.stderr or os.path.exists(str(tmp_path / "b.index") + ".csr")


def test_synthetic_counts_hook(stub, tmp_path):
    """--synthetic N,D,density (benchmark hook, not in the reference): no --mtx; every rank generates ITS shard of the same
    matrix in memory (a cell's row depends only on the seed and the cell id), uploads it and runs the usual schedule."""
    log = tmp_path / "calls.log"
    env = dict(os.environ, LD_LIBRARY_PATH=str(stub), MMV_STUB_LOG=str(log))
    cmd = [os.path.join(HOST, "bin", "nb_vae"), "--synthetic", "3000,300,0.05", "--out", str(tmp_path / "o"), "--batch_size", "256",
           "--max_epoch", "1", "--nboot", "1", "--seed", "3", "--skip_param_record", "--ngpu", "2"]
    p = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert p.returncode == 0, p.stderr[-2000:]
    ups = {}
    for ln in open(log):
        t = ln.split()
        if t[0] == "csr_upload":
            kv = dict(x.split("=", 1) for x in t[1:])
            ups[int(kv["dev"])] = (int(kv["cells"]), int(kv["genes"]), int(kv["nnz"]), kv["covar"])
    assert set(ups) == {0, 1}
    for dev in (0, 1):
        cells, genes, nnz, covar = ups[dev]
        assert (cells, genes, covar) == (1500, 300, "0")
        assert abs(nnz / (cells * genes) - 0.05) < 0.005          # geometric gaps: density * D nonzeros per cell on average
    assert not os.path.exists(str(tmp_path / "o") + ".covar.mtx.gz")   # nothing to mirror: no input file
