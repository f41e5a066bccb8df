#!/usr/bin/env python
"""Attributes the per-instruction counters of an `ncu --set full --import-source on` report to source
lines: executed warp instructions, stall samples and shared-memory wavefronts per line of the .cu file.

    python profiles/ncu_by_line.py REPORT.ncu-rep CUBIN MANGLED_KERNEL_NAME SOURCE.cu [units] [top]

CUBIN: `cuobjdump -xelf all libhcg.so` -> hcg_tc.sm_100a.cubin (built with -lineinfo).  Joins ncu's SASS page
with `nvdisasm -g` (line info per instruction) by code offset.
"""
import collections
import csv
import io
import re
import subprocess
import sys


def main():
    rep, cubin, kern, srcf = sys.argv[1:5]
    units = float(sys.argv[5]) if len(sys.argv) > 5 else 1.0
    top = int(sys.argv[6]) if len(sys.argv) > 6 else 50
    dis = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
    start = next(i for i, l in enumerate(dis) if l.startswith(".text." + kern + ":"))
    cur, line_of = None, {}
    for ln in dis[start + 1:]:
        if ln.startswith("\t.section") or ln.startswith(".text."):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
        if m:
            cur = (m.group(3) or m.group(1)).split("/")[-1], int(m.group(4) or m.group(2))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ix = {k: i for i, k in enumerate(hdr)}
    body = [r for r in rows[2:] if len(r) >= len(hdr)]
    base = int(body[0][0], 16)
    inst, samp, wave, ideal = (collections.Counter() for _ in range(4))
    for r in body:
        k = line_of.get(int(r[0], 16) - base)
        inst[k] += int(r[ix["Instructions Executed"]])
        samp[k] += int(r[ix["# Samples"]])
        wave[k] += int(r[ix["L1 Wavefronts Shared"]] or 0)
        ideal[k] += int(r[ix["L1 Wavefronts Shared Ideal"]] or 0)
    src = open(srcf).read().splitlines()
    name = srcf.split("/")[-1]
    ti, ts, tw = sum(inst.values()), sum(samp.values()), sum(wave.values())
    print(f"total: {ti / units:.1f} warp instructions per unit, {tw / units:.1f} shared wavefronts per unit "
          f"({sum(ideal.values()) / units:.1f} ideal)")
    print(f"{'inst/unit':>10s} {'inst%':>6s} {'samples%':>8s} {'wavefr/unit':>11s} {'ideal':>8s}  line")
    for k, n in sorted(inst.items(), key=lambda kv: -kv[1])[:top]:
        txt = src[k[1] - 1].strip()[:100] if k and k[0] == name and k[1] <= len(src) else ""
        print(f"{n / units:10.1f} {100 * n / ti:6.1f} {100 * samp[k] / max(ts, 1):8.1f} {wave[k] / units:11.1f} {ideal[k] / units:8.1f}  {k}  {txt}")


if __name__ == "__main__":
    main()
