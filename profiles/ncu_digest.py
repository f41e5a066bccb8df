#!/usr/bin/env python
"""Digest of one `ncu --set full --import-source on` report: headline metrics, stall reasons,
SASS opcode mix per unit of work and the most-sampled instructions.

    python profiles/ncu_digest.py gpurun_out/x.ncu-rep [units_per_launch] [top_n]
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
]


def page(rep, name, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    topn = int(sys.argv[3]) if len(sys.argv) > 3 else 12
    raw = page(rep, "raw")
    hdr, unit, val = raw[0], raw[1], raw[2]
    print("## metrics")
    for h, u, v in zip(hdr, unit, val):
        if h in KEYS or h == "Kernel Name":
            print(f"{h:75s} {v} {u}")
    print("## stalls per issue")
    st = [(h, float(v)) for h, v in zip(hdr, val) if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    for h, v in sorted(st, key=lambda t: -t[1])[:8]:
        print(f"  {h.split('issue_stalled_')[1].split('_per_issue')[0]:28s} {v:.2f}")
    rows = page(rep, "source", ["--print-source", "sass"])
    h = rows[1]
    ix = {k: i for i, k in enumerate(h)}
    body = [r for r in rows[2:] if len(r) >= len(h)]
    ops, samp = collections.Counter(), collections.Counter()
    tot = tots = 0
    for r in body:
        m = re.match(r"\s*(@!?U?P\w+\s+)?([A-Z0-9_]+)", r[ix["Source"]])
        op = m.group(2) if m else "?"
        n, s = int(r[ix["Instructions Executed"]]), int(r[ix["# Samples"]])
        ops[op] += n; samp[op] += s; tot += n; tots += s
    print(f"## SASS mix: {tot} warp instructions, {tot / units:.1f} per unit")
    for op, n in ops.most_common(22):
        print(f"  {op:10s} {n / units:9.1f}/unit {100 * n / tot:5.1f}%   samples {100 * samp[op] / max(tots, 1):5.1f}%")
    print("## most-sampled instructions")
    for i in sorted(range(len(body)), key=lambda i: -int(body[i][ix["# Samples"]]))[:topn]:
        r = body[i]
        print(f"--- {100 * int(r[ix['# Samples']]) / max(tots, 1):.1f}% of samples")
        for j in range(max(0, i - 4), min(len(body), i + 2)):
            q = body[j]
            print(f"   {'>' if j == i else ' '} {q[ix['Source']][:84]:84s} ex={q[ix['Instructions Executed']]} s={q[ix['# Samples']]}")


if __name__ == "__main__":
    main()
