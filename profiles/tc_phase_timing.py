#!/usr/bin/env python
"""Per-phase cycle breakdown of k_tc_fwd (thread 0 of CTA 0, clock64 deltas summed over its graphs).

Build the instrumented library here (needs nvcc), run on the GPU box:

    python profiles/tc_phase_timing.py --build          # -> build/libhcg_timing.so
    HCG_LIB_PATH=build/libhcg_timing.so python profiles/tc_phase_timing.py

Development tool; the product library is built without HCG_TC_TIMING.
"""
import ctypes as C
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "build", "libhcg_timing.so")
PHASES = ["loop top", "wait input stage", "prefetch issue + row setup + (row, norm) records", "slot barrier",
          "layer-0 gather + split + tcgen05.st", "tcgen05.wait::st + A-operand barrier (+ MMA issue)", "wait MMA",
          "epilogue (tcgen05.ld, bias, lrelu, sts)", "slot barrier", "layer-1 gather + split + tcgen05.st",
          "pooling + emb store"]

if "--build" in sys.argv:
    from hcatgnet_b200 import build as b
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = [b.find_nvcc(), *[f for f in b.NVCC_FLAGS if f != "-Xptxas=-v"], "-DHCG_TC_TIMING", "-shared", "-o", OUT, *b.sources()]
    subprocess.run(cmd, check=True)
    print("built", OUT)
    sys.exit(0)

import torch  # noqa: E402

import hcatgnet_b200 as pkg  # noqa: E402
from hcatgnet_b200 import _lib, ops  # noqa: E402
from hcatgnet_b200.synth import synth_batch  # noqa: E402

lib = _lib.load()
lib.hcg_debug_tc_timing.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
dev = torch.device("cuda", 0)
model = pkg.make_network("GCN", pkg.default_options(), 25).to(dev).eval()
dims, params = model._dims(), [p.detach() for p in model._flat_params()]
b = synth_batch(16384, 25, seed=1).to(dev)
plan = ops.collate(b.edge_index, b.batch, b.num_nodes, b.num_graphs, validate=False)
with torch.no_grad():
    for _ in range(3):
        ops.gcn_apply(b.x, plan, dims, params)
    torch.cuda.synchronize()
    lib.hcg_debug_tc_timing(None, 1)
    ops.gcn_apply(b.x, plan, dims, params)
    torch.cuda.synchronize()
buf = (C.c_ulonglong * 16)()
lib.hcg_debug_tc_timing(buf, 0)
graphs = 16384 / (148 * 2)
tot = sum(buf[:len(PHASES)])
print(f"thread 0 / CTA 0: ~{graphs:.1f} graphs, {tot / graphs:.0f} cycles per graph")
for i, name in enumerate(PHASES):
    print(f"  {i:2d} {name:42s} {buf[i] / graphs:8.0f} cyc/graph  {100 * buf[i] / tot:5.1f}%")
