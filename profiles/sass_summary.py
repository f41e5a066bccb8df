#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that show which hardware path a kernel uses (B200_PROFILING.md,
"What proves a Blackwell-native kernel"): UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UBLKCP = cp.async.bulk,
LDGSTS = cp.async, SYNCS = mbarrier, HMMA = legacy mma.sync.

    python profiles/sass_summary.py [hcatgnet_b200/libhcg.so] > profiles/r2_sass_summary.txt
"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "hcatgnet_b200/libhcg.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
PAT = [("UTC*MMA", r"\bUTC\w*MMA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UBLKCP", r"\bUBLKCP"), ("LDGSTS", r"\bLDGSTS"),
       ("SYNCS", r"\bSYNCS"), ("HMMA", r"\bHMMA"), ("FFMA", r"\bFFMA"), ("LDS", r"\bLDS"), ("ATOMS", r"\bATOMS"), ("BAR", r"\bBAR\.")]
cur, counts, total = None, collections.OrderedDict(), collections.Counter()
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"^void ", "", cur).replace("hcg::", "")
        cur = re.sub(r"\(.*$", "", cur)
        counts[cur] = collections.Counter()
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
        total[cur] += 1
        for name, pat in PAT:
            if re.search(pat, ln):
                counts[cur][name] += 1
print(f"{'kernel':52s} {'instr':>6s} " + " ".join(f"{n:>7s}" for n, _ in PAT))
for k, c in counts.items():
    print(f"{k[:52]:52s} {total[k]:6d} " + " ".join(f"{c[n]:7d}" for n, _ in PAT))
