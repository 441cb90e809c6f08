#!/usr/bin/env python
"""Golden vectors for the bgzf Matrix-Market reader, produced by the REFERENCE
reader itself (oracle/_ref/ref_reader_harness).  Build container only.

Writes into tests/golden/:
  reader_a.mtx.gz   many tiny bgzf blocks, comments, gaps, duplicates, quirky lines
  reader_b.mtx.gz   integer counts, 64 KiB blocks (as bgzip would write)
  reader_ones.mtx.gz  the reference's own create_ones_like output (its bgzf writer)
  reader_golden.npz   index text, index tables, headers, block triplets, dense
                      minibatches, take_float results — all from the reference.
"""
import gzip
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
READER = os.path.join(ROOT, "oracle", "_ref", "ref_reader_harness")
sys.path.insert(0, ROOT)

FLOAT_TOKENS = ("0 1 2 10 57 123456789 0.5 .5 1.5 3.14159274 2e-3 1E2 1e+2 -3.25 007 12345678.9 "
                "0.000001 1.0e-10 9.999999e5 16777217 0.1 0.2 0.3 1e0 5. -0 1.17549435e-38 3e38 "
                "4294967296 1.23456789012345 00.50 7e-1 2.5E+3 1e10 1e-5 33.333333 99999.9").split()


def make_file_a(rng):
    """D=40 features x N=60 cells, float values, irregular text."""
    D, N = 40, 60
    lines = ["%%MatrixMarket matrix coordinate real general", "% a comment", "%another"]
    body = []
    empty_cols = {5, 6, 7, 30, 58}           # interior empty columns (never the last)
    vals = ["1", "2", "3.5", "1e1", "0.25", "7", "12", "2.0E-1", "100", "4"]
    nnz = 0
    for col in range(N):
        if col in empty_cols:
            continue
        rows = sorted(rng.choice(D, size=rng.integers(1, 9), replace=False))
        for r in rows:
            body.append(f"{r + 1} {col + 1} {vals[rng.integers(len(vals))]}")
            nnz += 1
        if col == 10:
            body.append(f"{rows[0] + 1} {col + 1} 9")      # duplicate (row, col): last wins
            nnz += 1
        if col == 12:
            body.append("% comment in the middle")
            body.append("3 13")                             # incomplete line -> skipped
        if col == 20:
            body.append(f"\t{rows[0] + 1}\t{col + 1}   8.5  ")  # tabs / repeated blanks
            nnz += 1
        if col == 21:
            body.append("")                                 # empty line
    lines.append(f"{D} {N} {nnz}")
    text = "\n".join(lines + body) + "\n"
    return text.encode(), D, N


def make_file_b(rng):
    """D=300 x N=500 integer counts, ~8% dense -> several 64 KiB blocks."""
    D, N = 300, 500
    body = []
    for col in range(N):
        k = max(1, rng.binomial(D, 0.08))
        rows = sorted(rng.choice(D, size=k, replace=False))
        for r in rows:
            body.append(f"{r + 1} {col + 1} {1 + rng.poisson(1.2)}")
    head = ["%%MatrixMarket matrix coordinate integer general", f"{D} {N} {len(body)}"]
    return ("\n".join(head + body)).encode(), D, N   # NO trailing newline at EOF


def run(*args):
    subprocess.run([READER, *args], check=True, stderr=subprocess.DEVNULL,
                   stdout=subprocess.PIPE)


def batches_for(N, B, rng):
    """Sequential wrap-around batches (mmvae_alg.hh:264-266) + adversarial ones."""
    out = []
    nb = (N + B - 1) // B
    for b in (0, 1, nb - 1):
        out.append([(b * B + j) % N for j in range(B)])
    out.append(list(rng.integers(0, N, B)))                    # random, with repeats
    out.append([3] * B)                                         # all the same cell
    out.append([N - 1] + [0] * (B - 1))                         # last column reads to EOF
    out.append(sorted(rng.choice(N, B, replace=False))[::-1])   # descending
    return np.array(out, dtype=np.int64)


def main():
    from oracle import bgzf_mtx
    rng = np.random.default_rng(20201)
    out = {}
    tmp = tempfile.mkdtemp(prefix="rgold_")

    files = {}
    data, D, N = make_file_a(rng)
    bgzf_mtx.bgzf_write(os.path.join(HERE, "reader_a.mtx.gz"), data, block_size=257)
    files["a"] = (os.path.join(HERE, "reader_a.mtx.gz"), D, N, 7)
    data, D, N = make_file_b(rng)
    bgzf_mtx.bgzf_write(os.path.join(HERE, "reader_b.mtx.gz"), data)
    files["b"] = (os.path.join(HERE, "reader_b.mtx.gz"), D, N, 32)
    ones = os.path.join(HERE, "reader_ones.mtx.gz")
    run("ones", "500", ones)
    files["ones"] = (ones, 1, 500, 32)

    for tag, (path, D, N, B) in files.items():
        idx = os.path.join(tmp, tag + ".index")
        run("index", path, idx)
        out[f"{tag}__index_text"] = np.frombuffer(gzip.open(idx, "rb").read(), dtype=np.uint8)
        run("table", idx, os.path.join(tmp, tag + ".tab"))
        tab = np.fromfile(os.path.join(tmp, tag + ".tab"), dtype=np.int64)
        out[f"{tag}__table"] = tab
        hdr = subprocess.run([READER, "header", path], check=True, stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL).stdout.split()
        out[f"{tag}__header"] = np.array([int(v) for v in hdr], dtype=np.int64)
        # whole-file triplets and a middle block
        for name, (beg, end) in {"all": (int(tab[0]), 0),
                                 "mid": (int(tab[N // 3]), int(tab[N // 3 + 4]))}.items():
            pre = os.path.join(tmp, f"{tag}_{name}")
            run("block", path, str(beg), str(end), pre)
            out[f"{tag}__block_{name}_range"] = np.array([beg, end], dtype=np.int64)
            out[f"{tag}__block_{name}_rows"] = np.fromfile(pre + ".rows.i64", dtype=np.int64)
            out[f"{tag}__block_{name}_cols"] = np.fromfile(pre + ".cols.i64", dtype=np.int64)
            out[f"{tag}__block_{name}_vals"] = np.fromfile(pre + ".vals.f32", dtype=np.float32)
        bt = batches_for(N, B, rng)
        bt.tofile(os.path.join(tmp, tag + ".cols"))
        run("batch", path, idx, os.path.join(tmp, tag + ".cols"), str(B), os.path.join(tmp, tag + ".dense"))
        out[f"{tag}__batches"] = bt
        out[f"{tag}__dense"] = np.fromfile(os.path.join(tmp, tag + ".dense"),
                                           dtype=np.float32).reshape(len(bt), B, D)

    with open(os.path.join(tmp, "floats.txt"), "w") as f:
        f.write("\n".join(FLOAT_TOKENS) + "\n")
    run("floats", os.path.join(tmp, "floats.txt"), os.path.join(tmp, "floats.f32"))
    out["float_tokens"] = np.frombuffer("\n".join(FLOAT_TOKENS).encode(), dtype=np.uint8)
    out["float_values"] = np.fromfile(os.path.join(tmp, "floats.f32"), dtype=np.float32)

    np.savez_compressed(os.path.join(HERE, "reader_golden.npz"), **out)
    print("reader_golden.npz:", len(out), "arrays,",
          os.path.getsize(os.path.join(HERE, "reader_golden.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
