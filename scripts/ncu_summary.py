#!/usr/bin/env python
"""Summaries of an `ncu --set full` report for profiles/ (usage: ncu_summary.py <report.ncu-rep> <out prefix> [kernel regex for
the SASS breakdown]): <prefix>_summary.csv = one row per captured launch with the counters DESIGN.md quotes;
<prefix>_sass_<kernel>.txt = the kernel's executed-instruction histogram by opcode and its hottest SASS lines."""
import collections
import csv
import io
import re
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "lts__t_bytes.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "launch__shared_mem_per_block_dynamic", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
           "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, check=True).stdout


def main():
    rep, prefix = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
    h, units = rows[0], rows[1]
    cols = [m for m in METRICS if m in h]
    with open(prefix + "_summary.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [f"{m} [{units[h.index(m)]}]" for m in cols])
        for r in rows[2:]:
            w.writerow([r[h.index("Kernel Name")][:90]] + [r[h.index(m)] for m in cols])
    if len(sys.argv) < 4:
        return
    pat = sys.argv[3]
    src = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat, "--print-source", "sass"))))
    hh = src[1]
    isrc, ie, iss = hh.index("Source"), hh.index("Instructions Executed"), hh.index("Warp Stall Sampling (All Samples)")
    data = [(r[isrc].strip(), int(r[ie]), int(r[iss])) for r in src[2:] if len(r) > ie and r[ie].isdigit()]
    tot, tots = sum(d[1] for d in data), sum(d[2] for d in data)
    op, ops = collections.Counter(), collections.Counter()
    for s, e, st in data:
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", s)
        o = m.group(2).split(".")[0] if m else "?"
        op[o] += e
        ops[o] += st
    with open(f"{prefix}_sass_{re.sub('[^a-z0-9_]', '', pat.lower())}.txt", "w") as f:
        f.write(f"# {src[0][1] if len(src[0]) > 1 else pat}\n# warp instructions executed {tot}, stall samples {tots}, SASS lines {len(data)}\n")
        f.write("# opcode, executed, share of executed, share of stall samples\n")
        for o, c in op.most_common(28):
            f.write(f"{o:12s} {c:12d} {100 * c / tot:5.1f}% {100 * ops[o] / max(tots, 1):5.1f}%\n")
        f.write("# hottest SASS lines by executed count x: index, executed, stall samples, instruction\n")
        for i, (s, e, st) in sorted(enumerate(data), key=lambda t: -t[1][1])[:25]:
            f.write(f"{i:5d} {e:10d} {st:6d}  {s[:110]}\n")
        f.write("# lines with the most stall samples\n")
        for i, (s, e, st) in sorted(enumerate(data), key=lambda t: -t[1][2])[:25]:
            f.write(f"{i:5d} {e:10d} {st:6d}  {s[:110]}\n")


if __name__ == "__main__":
    main()
