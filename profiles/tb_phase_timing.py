#!/usr/bin/env python
"""Per-phase cycle breakdown of k_tc_bwd (thread 0 of CTA 0, clock64 deltas summed over its tiles).

    python profiles/tc_phase_timing.py --build          # -> build/libhcg_timing.so (same instrumented build)
    HCG_LIB_PATH=build/libhcg_timing.so python profiles/tb_phase_timing.py
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import hcatgnet_b200 as pkg  # noqa: E402
from hcatgnet_b200 import _lib  # noqa: E402
from hcatgnet_b200.synth import synth_batch  # noqa: E402

PHASES = ["loop top", "wait CSR stage", "issue next + h2 load + pooling data", "barrier", "records + P1 ties + barrier + P1 apply",
          "wait previous P7 (g_W1) + readback", "h1 load (raw + residual)", "barrier", "P2 gather + Q tiles + tcgen05.st",
          "fences + barrier + MMA issue + wait P4", "P4 epilogue", "barrier + P5 column sums", "wait P3 (g_W2) + readback",
          "x load (raw + residual)", "P6 gather + Q tiles + barrier + P7 issue"]
lib = _lib.load()
lib.hcg_debug_tb_timing.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
model = pkg.make_network("GCN", pkg.default_options(), 25).to("cuda").train()
tr = pkg.FusedTrainer(model)
b = synth_batch(4096, 25, seed=1).to("cuda")
for _ in range(3):
    tr.step(b)
torch.cuda.synchronize()
lib.hcg_debug_tb_timing(None, 1)
tr.step(b)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 16)()
lib.hcg_debug_tb_timing(buf, 0)
tot = sum(buf[:len(PHASES)])
graphs = 4096 / 148
print(f"thread 0 / CTA 0: ~{graphs:.1f} graphs, {tot / graphs:.0f} cycles per graph, {tot} cycles in all")
for i, name in enumerate(PHASES):
    print(f"  {i:2d} {name:52s} {buf[i] / graphs:8.0f} cyc/graph  {100 * buf[i] / tot:5.1f}%")
