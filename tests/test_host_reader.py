"""Host C++ surface (mm-vae_b200/host): reader + option parsers, CPU only.

The tool's output is compared bit for bit with tests/golden/reader_golden.npz, i.e. with what the
REFERENCE reader (mmutil_index.hh, mmutil_bgzf_util.hh, strbuf.hh, bgzf.c compiled verbatim by
oracle/_ref/ref_reader_harness) produced on the same files — the same fixtures the Python oracle is
pinned to in test_oracle_golden.py.
"""
import gzip
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
HOST = os.path.join(ROOT, "mm-vae_b200", "host")
TOOL = os.path.join(HOST, "bin", "mmv_reader_tool")
FILES = {"a": "reader_a.mtx.gz", "b": "reader_b.mtx.gz", "ones": "reader_ones.mtx.gz"}


@pytest.fixture(scope="module")
def tool():
    subprocess.run(["make", "-C", HOST, "bin/mmv_reader_tool"], check=True, stdout=subprocess.DEVNULL)
    return TOOL


@pytest.fixture(scope="module")
def rg():
    z = np.load(os.path.join(GOLDEN, "reader_golden.npz"))
    return {k: z[k] for k in z.files}


def run(tool, *args, env=None):
    return subprocess.run([tool, *map(str, args)], check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                          text=True, env=None if env is None else dict(os.environ, **env)).stdout


def test_take_float_bit_exact(tool, rg):
    toks = bytes(rg["float_tokens"]).decode().split("\n")
    got = [int(v) for v in run(tool, "floats", *toks).split()]
    assert got == rg["float_values"].view(np.uint32).tolist()


@pytest.mark.parametrize("tag", list(FILES))
def test_header_index_table(tool, rg, tag, tmp_path):
    path = os.path.join(GOLDEN, FILES[tag])
    out = run(tool, "header", path).split("\n")
    assert [int(v) for v in out[0].split()] == rg[f"{tag}__header"].tolist()
    assert out[1] == "is_bgzf 1"
    assert run(tool, "scan", path).encode() == bytes(rg[f"{tag}__index_text"])
    idx = tmp_path / "t.index"
    tab = [int(v) for v in run(tool, "table", path, idx).split()]
    assert gzip.open(idx, "rb").read() == bytes(rg[f"{tag}__index_text"])   # same index FILE format
    assert tab == rg[f"{tag}__table"].tolist()


@pytest.mark.parametrize("tag", list(FILES))
@pytest.mark.parametrize("blk", ["all", "mid"])
def test_block_triplets(tool, rg, tag, blk):
    path = os.path.join(GOLDEN, FILES[tag])
    beg, end = rg[f"{tag}__block_{blk}_range"].tolist()
    rows = np.array([[int(v) for v in ln.split()] for ln in run(tool, "block", path, beg, end).splitlines()], dtype=np.int64)
    assert rows[:, 0].tolist() == rg[f"{tag}__block_{blk}_rows"].tolist()
    assert rows[:, 1].tolist() == rg[f"{tag}__block_{blk}_cols"].tolist()
    assert rows[:, 2].tolist() == rg[f"{tag}__block_{blk}_vals"].view(np.uint32).tolist()


@pytest.mark.parametrize("tag", list(FILES))
def test_dense_minibatches_indexed_and_csr(tool, rg, tag, tmp_path):
    """Minibatch indexing is bit-exact both through the reference's index/block route and through the
    parse-once CSR the device path uses (wrap-around batches, duplicate cells included)."""
    path = os.path.join(GOLDEN, FILES[tag])
    idx = tmp_path / "t.index"
    run(tool, "table", path, idx)
    D = int(rg[f"{tag}__header"][0])
    for cols, dense in zip(rg[f"{tag}__batches"], rg[f"{tag}__dense"]):
        out = tmp_path / "d.bin"
        run(tool, "dense", path, idx, out, *cols.tolist())
        got = np.fromfile(out, dtype=np.float32).reshape(2, len(cols), D)
        assert np.array_equal(got[0].view(np.uint32), dense.view(np.uint32))
        assert np.array_equal(got[1].view(np.uint32), dense.view(np.uint32))


def test_csr_rows_sorted_and_complete(tool, rg, tmp_path):
    path = os.path.join(GOLDEN, FILES["b"])
    idx = tmp_path / "t.index"
    run(tool, "table", path, idx)
    D, N, nnz = [int(v) for v in run(tool, "csr", path, idx, tmp_path / "c").split()]
    indptr = np.fromfile(str(tmp_path / "c") + ".indptr", dtype=np.int64)
    indices = np.fromfile(str(tmp_path / "c") + ".indices", dtype=np.int32)
    values = np.fromfile(str(tmp_path / "c") + ".values", dtype=np.float32)
    assert [D, N] == rg["b__header"][:2].tolist() and indptr.shape == (N + 1,) and indptr[-1] == nnz == len(indices)
    for c in range(N):
        seg = indices[indptr[c]:indptr[c + 1]]
        assert np.all(np.diff(seg) > 0) and (len(seg) == 0 or (seg[0] >= 0 and seg[-1] < D))
    # every triplet of the full-file block is in the CSR (last duplicate wins)
    last = {}
    for r, c, v in zip(rg["b__block_all_rows"], rg["b__block_all_cols"], rg["b__block_all_vals"]):
        last[(int(c), int(r))] = v
    assert len(last) == nnz
    for (c, r), v in list(last.items())[::37]:
        seg = indices[indptr[c]:indptr[c + 1]]
        k = int(np.searchsorted(seg, r))
        assert seg[k] == r and values[indptr[c] + k].view(np.uint32) == np.float32(v).view(np.uint32)


def test_bgzf_writer_roundtrip(tool, rg, tmp_path):
    txt = tmp_path / "m.mtx"
    lines = ["%%MatrixMarket matrix coordinate integer general", "3 5 5", "1 1 2", "3 1 1", "2 3 7", "1 4 1.5e1", "3 5 4"]
    txt.write_text("\n".join(lines) + "\n")
    gz = tmp_path / "m.mtx.gz"
    run(tool, "write", txt, gz)
    assert gzip.open(gz, "rb").read().decode() == txt.read_text()           # valid multi-member gzip
    assert run(tool, "header", gz).split("\n")[0] == "3 5 5"
    # column 2 (0-based 1) is empty: the gap fill gives it the next column's offset (mmutil_index.hh:219-224)
    tab = [int(v) for v in run(tool, "table", gz, tmp_path / "m.index").split()]
    assert len(tab) == 5 and tab[1] == tab[2] and tab[0] < tab[2] < tab[3] < tab[4]


def test_option_defaults_and_aliases(tool, tmp_path):
    mtx = os.path.join(GOLDEN, FILES["a"])
    kv = dict(ln.split(" ", 1) if " " in ln else (ln, "") for ln in run(tool, "opts", "--mtx", mtx, "--out", "o").splitlines())
    # defaults: mmvae.hh:36-39, mmvae_alg.hh:19-23, nb.hh:58-61, vmf.hh:59-62
    assert kv["rc"] == "0" and kv["idx"] == mtx + ".index" and kv["batch_size"] == "100"
    assert (kv["kl_discount"], kv["kl_min"], kv["kl_max"]) == ("0.1", "0.01", "1")
    assert (kv["lr"], kv["grad_clip"], kv["nboot"], kv["max_epoch"], kv["recording"]) == ("0.001", "1", "3", "101", "10")
    assert (kv["mean_latent"], kv["overdispersion_encoding"], kv["overdispersion_latent"], kv["nb_relu"]) == ("2", "1", "1", "0")
    assert (kv["latent"], kv["kappa_min"], kv["kappa_max"], kv["vmf_relu"]) == ("2", "0.1", "10", "0")
    assert kv["vmf_latent"] == "0" and kv["csr_cache"] == "0"   # options the reference does not have: off by default
    kv = dict(ln.split(" ", 1) if " " in ln else (ln, "") for ln in run(
        tool, "opts", "--mtx", mtx, "--output", "o", "--batch", "64", "--cov", "c.mtx.gz", "--kl_min", "0.5",
        "--learning_rate", "0.01", "--bootstrap", "5", "--epoch", "7", "--mean-encoding", "128,64", "--mean_decoding", "32",
        "--mean_latent", "10", "--overdisp-encoding", "3", "--overdispersion-latent", "2", "--relu", "--encoding", "16,8",
        "--latent", "4", "--kappa-max", "20", "--seed", "9", "--dtype", "bf16", "--train_all_layers", "--vmf_latent", "--csr_cache").splitlines())
    assert kv["out"] == "o" and kv["batch_size"] == "64" and kv["covar_mtx"] == "c.mtx.gz" and kv["covar_idx"] == "c.mtx.gz.index"
    assert kv["kl_min"] == "0.5" and kv["lr"] == "0.01" and kv["nboot"] == "5" and kv["max_epoch"] == "7"
    assert kv["mean_encoding"] == "128,64" and kv["mean_decoding"] == "32" and kv["mean_latent"] == "10"
    assert kv["overdispersion_encoding"] == "3" and kv["overdispersion_latent"] == "2" and kv["nb_relu"] == "1"
    assert kv["encoding"] == "16,8" and kv["latent"] == "4" and kv["kappa_max"] == "20" and kv["vmf_relu"] == "1"
    assert kv["seed"] == "9" and kv["bf16"] == "1" and kv["train_all_layers"] == "1" and kv["vmf_latent"] == "1"
    assert kv["csr_cache"] == "1"
    # missing --out fails like the reference (ERR_RET "need output file header")
    p = subprocess.run([tool, "opts", "--mtx", mtx], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert "rc 1" in p.stdout and "need output file header" in p.stderr


def test_covar_ones_matrix_writer(tool, tmp_path):
    """No --covar: the reference writes ${out}.covar.mtx.gz, a 1 x N matrix of ones as bgzf Matrix Market text
    sorted by column, and indexes it (nb_vae_main.cc:67-73, mmvae_io.hh:292-310, io.hh:191-226).  The file
    must read back through the same reader: N columns, each holding the single triplet (row 0, value 1)."""
    N = 70000   # more than one 64 KB bgzf block of text
    path = tmp_path / "run.covar.mtx.gz"
    run(tool, "ones", N, path)
    text = gzip.open(path, "rt").read().split("\n")
    assert text[0] == "%%MatrixMarket matrix coordinate integer general"
    assert text[1] == f"1 {N} {N}"
    assert text[2] == "1 1 1" and text[N + 1] == f"1 {N} 1" and text[N + 2] == ""
    assert [int(v) for v in run(tool, "header", path).split("\n")[0].split()] == [1, N, N]
    idx = [l.split() for l in gzip.open(str(path) + ".index", "rt").read().strip().split("\n")]
    assert len(idx) == N and [int(r[0]) for r in idx] == list(range(N))
    prefix = tmp_path / "ones"
    dims = [int(v) for v in run(tool, "csr", path, str(path) + ".index", prefix).split()]
    assert dims == [1, N, N]
    assert np.array_equal(np.fromfile(str(prefix) + ".indptr", np.int64), np.arange(N + 1))
    assert not np.fromfile(str(prefix) + ".indices", np.int32).any()
    assert np.all(np.fromfile(str(prefix) + ".values", np.float32) == 1.0)


@pytest.mark.parametrize("tag", list(FILES))
def test_parallel_ingest_equals_sequential(tool, tag, tmp_path):
    """The one-off ingest splits the text at column boundaries of the index table and parses the ranges on
    several host threads; merging in file order must give byte-identical CSR arrays to the 1-thread read
    (duplicates: last one wins; unsorted rows inside a column: sorted by gene)."""
    path = os.path.join(GOLDEN, FILES[tag])
    idx = tmp_path / "t.index"
    run(tool, "table", path, idx)
    outs = []
    for threads in ("1", "3", "8"):
        prefix = tmp_path / f"csr{threads}"
        env = dict(os.environ, MMV_INGEST_THREADS=threads)
        subprocess.run([tool, "csr", path, str(idx), str(prefix)], check=True, stdout=subprocess.PIPE, env=env)
        outs.append([open(str(prefix) + ext, "rb").read() for ext in (".indptr", ".indices", ".values")])
    assert outs[0] == outs[1] == outs[2]
    assert len(outs[0][1]) > 0


def test_binary_csr_cache(tool, tmp_path):
    """--csr_cache (not in the reference): ${idx}.csr holds the parsed CSR; it is used only while the
    mtx file is byte-identical (size, mtime, index table), and gives byte-identical arrays."""
    import shutil
    import time
    src = os.path.join(GOLDEN, FILES["b"])
    mtx = tmp_path / "m.mtx.gz"
    shutil.copy(src, mtx)
    idx = tmp_path / "m.index"
    run(tool, "table", mtx, idx)
    arrays = lambda p: [np.fromfile(str(tmp_path / p) + s, dtype=np.uint8) for s in (".indptr", ".indices", ".values")]
    run(tool, "csr", mtx, idx, tmp_path / "plain")
    out = run(tool, "csr", mtx, idx, tmp_path / "first", "cache").split("\n")
    assert out[1] == "parsed" and os.path.exists(str(idx) + ".csr") and not os.path.exists(str(idx) + ".csr.tmp")
    out2 = run(tool, "csr", mtx, idx, tmp_path / "second", "cache").split("\n")
    assert out2[1] == "cache" and out2[0] == out[0]
    for a, b, c in zip(arrays("plain"), arrays("first"), arrays("second")):
        assert np.array_equal(a, b) and np.array_equal(a, c)
    # the environment variable turns it on without the option
    env = dict(os.environ, MMV_CSR_CACHE="1")
    p = subprocess.run([tool, "csr", str(mtx), str(idx), str(tmp_path / "env")], check=True, env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True)
    assert "binary CSR cache" in p.stderr
    # a touched file (same bytes, new mtime) and a different file both invalidate it
    time.sleep(0.01)
    os.utime(mtx, ns=(time.time_ns(), time.time_ns()))
    assert run(tool, "csr", mtx, idx, tmp_path / "third", "cache").split("\n")[1] == "parsed"
    assert run(tool, "csr", mtx, idx, tmp_path / "fourth", "cache").split("\n")[1] == "cache"
    shutil.copy(os.path.join(GOLDEN, FILES["a"]), mtx)
    run(tool, "table", mtx, idx)
    out5 = run(tool, "csr", mtx, idx, tmp_path / "fifth", "cache").split("\n")
    assert out5[1] == "parsed" and out5[0] != out[0]
    # a truncated cache is rejected, not trusted
    blob = open(str(idx) + ".csr", "rb").read()
    open(str(idx) + ".csr", "wb").write(blob[:-3])
    assert run(tool, "csr", mtx, idx, tmp_path / "sixth", "cache").split("\n")[1] == "parsed"
    for a, b in zip(arrays("fifth"), arrays("sixth")):
        assert np.array_equal(a, b)


def test_synthetic_shard_roundtrip_through_mtx(tool, tmp_path):
    """a subsample of the benchmark's synthetic NB counts (bench.gen_shard, cfg-3 feature count) written as
    bgzf Matrix Market, indexed and read back by the host reader gives the identical CSR — the device-generated
    shards and the file path index cells and genes the same way."""
    import importlib.util
    import torch
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    n, D = 400, 20_000
    indptr, indices, values, longest = bench.gen_shard(n, D, 0.05, 7, torch.device("cpu"), chunk=128)
    indptr, indices, values = indptr.numpy(), indices.numpy(), values.numpy()
    assert 0.04 < len(indices) / (n * D) < 0.06 and (np.diff(indptr) >= 1).all() and longest == np.diff(indptr).max()
    cell = np.repeat(np.arange(n), np.diff(indptr))
    txt = tmp_path / "s.mtx"
    with open(txt, "w") as f:
        f.write(f"%%MatrixMarket matrix coordinate integer general\n{D} {n} {len(indices)}\n")
        f.write("\n".join(f"{g + 1} {c + 1} {int(v)}" for g, c, v in zip(indices.tolist(), cell.tolist(), values.tolist())))
        f.write("\n")
    gz, idx = tmp_path / "s.mtx.gz", tmp_path / "s.index"
    run(tool, "write", txt, gz)
    run(tool, "table", gz, idx)
    Dn, Nn, nnz = [int(v) for v in run(tool, "csr", gz, idx, tmp_path / "c").split()]
    assert (Dn, Nn, nnz) == (D, n, len(indices))
    assert np.array_equal(np.fromfile(str(tmp_path / "c") + ".indptr", dtype=np.int64), indptr)
    assert np.array_equal(np.fromfile(str(tmp_path / "c") + ".indices", dtype=np.int32), indices)
    assert np.array_equal(np.fromfile(str(tmp_path / "c") + ".values", dtype=np.float32), values)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_shard_ingest_concatenates_to_the_full_csr(tool, tag, rg, tmp_path):
    """mtx_data_block_t(..., col_lo, col_hi): every shard parses only its own range of the text; the shards'
    CSR arrays laid end to end are the whole-file CSR, for any cut (incl. cuts next to empty data points)."""
    gz = os.path.join(GOLDEN, FILES[tag])
    idx = tmp_path / "i.index"
    run(tool, "table", gz, idx)
    run(tool, "csr", gz, idx, tmp_path / "full", env={"MMV_INGEST_THREADS": "1"})
    full = {k: np.fromfile(str(tmp_path / "full") + "." + k, dtype=d) for k, d in (("indptr", np.int64), ("indices", np.int32), ("values", np.float32))}
    N = len(full["indptr"]) - 1
    for cuts in ([0, N], [0, N // 2, N], [0, 1, 2, N - 1, N], [0, N // 3, N // 3 + 1, 2 * N // 3, N]):
        cuts = sorted(set(cuts))
        ip, ix, vv = [np.zeros(1, np.int64)], [], []
        for th, (lo, hi) in enumerate(zip(cuts[:-1], cuts[1:])):
            out = run(tool, "shard", gz, idx, tmp_path / "s", lo, hi, env={"MMV_INGEST_THREADS": str(1 + th % 3)}).split()
            assert int(out[2]) == hi - lo
            p = np.fromfile(str(tmp_path / "s") + ".indptr", dtype=np.int64)
            assert len(p) == hi - lo + 1 and p[0] == 0
            ip.append(p[1:] + ip[-1][-1])
            ix.append(np.fromfile(str(tmp_path / "s") + ".indices", dtype=np.int32))
            vv.append(np.fromfile(str(tmp_path / "s") + ".values", dtype=np.float32))
        assert np.array_equal(np.concatenate(ip), full["indptr"]), cuts
        assert np.array_equal(np.concatenate(ix), full["indices"]), cuts
        assert np.array_equal(np.concatenate(vv).view(np.uint32), full["values"].view(np.uint32)), cuts


def test_reader_under_sanitizers(rg, tmp_path):
    """The host reader (bgzf inflate, in-place line split, index, parallel ingest, CSR cache, dense views) built with
    AddressSanitizer + UBSan: any finding aborts the tool and fails the call."""
    if subprocess.run(["make", "-C", HOST, "bin/mmv_reader_tool_asan"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode != 0:
        pytest.skip("no sanitizer runtime for the host compiler on this machine")
    asan = os.path.join(HOST, "bin", "mmv_reader_tool_asan")
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0", UBSAN_OPTIONS="halt_on_error=1")

    def go(*args, **extra):
        return subprocess.run([asan, *map(str, args)], check=True, env=dict(env, **extra), stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True).stdout
    for tag in ("a", "b"):
        gz = os.path.join(GOLDEN, FILES[tag])
        idx = tmp_path / f"{tag}.index"
        go("table", gz, idx)
        go("block", gz, 0, 0)
        for th in ("1", "5"):
            go("csr", gz, idx, tmp_path / f"c{tag}{th}", MMV_INGEST_THREADS=th)
        go("csr", gz, idx, tmp_path / "cc", "cache")
        assert go("csr", gz, idx, tmp_path / "cc", "cache").split("\n")[1] == "cache"
        n = int(rg[f"{tag}__header"][1])
        go("dense", gz, idx, tmp_path / "d.bin", 0, n - 1, n // 2, 0)
    go("floats", "1.5e1", "-3.25", "7", "1e-3", "00012.5000", ".5", "abc")
    go("ones", 37, tmp_path / "ones.mtx.gz")
    # corrupt / foreign blocks must be refused without touching memory outside the buffers: BSIZE below the
    # 26-byte minimum (header + trailer; 'bsize - 18 - 8' used to underflow), a truncated block, a wrong ISIZE
    raw = open(os.path.join(GOLDEN, FILES["a"]), "rb").read()
    bsize = raw[16] | (raw[17] << 8)
    cases = {"short_bsize": raw[:16] + bytes([20, 0]) + raw[18:],
             "truncated": raw[:bsize // 2],
             "bad_isize": raw[:bsize - 3] + bytes([0xff, 0xff, 0x7f, 0x00]) + raw[bsize + 1:]}
    for name, blob in cases.items():
        bad = tmp_path / f"{name}.mtx.gz"
        bad.write_bytes(blob)
        p = subprocess.run([asan, "block", str(bad), "0", "0"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert "AddressSanitizer" not in p.stderr and "runtime error" not in p.stderr, (name, p.stderr[-1500:])
        p = subprocess.run([asan, "table", str(bad), str(tmp_path / f"{name}.index")], env=env, stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, text=True)
        assert "AddressSanitizer" not in p.stderr and "runtime error" not in p.stderr, (name, p.stderr[-1500:])


def test_parallel_ingest_under_thread_sanitizer(tmp_path):
    """The multi-threaded one-off ingest (ranges parsed by separate readers, parallel tidy / CSR assembly) under TSan."""
    if subprocess.run(["make", "-C", HOST, "bin/mmv_reader_tool_tsan"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode != 0:
        pytest.skip("no TSan runtime for the host compiler on this machine")
    tsan = os.path.join(HOST, "bin", "mmv_reader_tool_tsan")
    gz = os.path.join(GOLDEN, FILES["b"])
    env = dict(os.environ, MMV_INGEST_THREADS="6", TSAN_OPTIONS="halt_on_error=1 exitcode=66")
    subprocess.run([tsan, "table", gz, str(tmp_path / "b.index")], check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    p = subprocess.run([tsan, "csr", gz, str(tmp_path / "b.index"), str(tmp_path / "c")], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True)
    assert p.returncode == 0 and "ThreadSanitizer" not in p.stderr, p.stderr[-2000:]
    assert "6 reader threads" in p.stderr or "reader threads" in p.stderr
