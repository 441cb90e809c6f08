#!/usr/bin/env python
"""Refresh profiles/traffic.json from an `ncu --set full` capture (usage: ncu_traffic.py <report.ncu-rep> <capture tag>):
per-launch dram__bytes_read.sum + dram__bytes_write.sum of the kernels bench.py reports a roofline for, stamped with the
hash of the kernel's CURRENT source (bench.py drops an entry whose source has changed since) and the current commit."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (src_sha, KERNEL_SOURCES)

PATTERNS = {"nb_likelihood": "nb_likelihood_fast_kernel", "vmf_likelihood": "vmf_likelihood_fast_kernel",
            "senc_fwd": "senc_tc_kernel<256, 0,", "senc_bwd": "senc_tc_kernel<256, 1,", "panel_build": "panel_build_kernel",
            "gemm_dec_last_fwd": "gemm_tc_persist_kernel<0, 0>", "gemm_dec_last_dw": "gemm_tc_persist_kernel<1, 1>",
            "clip_adam": "adam_kernel"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep, tag = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units = rows[0], rows[1]
    kn, rd, wr = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], stdout=subprocess.PIPE, text=True).stdout.strip()
    path = os.path.join(ROOT, "profiles", "traffic.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    for key, pat in PATTERNS.items():
        vals = [float(r[rd].replace(",", "")) * UNIT[units[rd]] + float(r[wr].replace(",", "")) * UNIT[units[wr]]
                for r in rows[2:] if pat in r[kn]]
        if not vals:
            continue
        out[key] = {"dram_bytes": sum(vals) / len(vals), "src_sha": bench.src_sha(*bench.KERNEL_SOURCES[key]),
                    "commit": commit, "capture": tag, "launches_averaged": len(vals)}
        print(key, out[key])
    json.dump(out, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
