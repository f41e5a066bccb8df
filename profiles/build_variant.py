#!/usr/bin/env python
"""Builds a variant of libhcg.so with extra nvcc flags (experiments; the product library is hcatgnet_b200/libhcg.so):

    python profiles/build_variant.py build/libhcg_x.so -DHCG_SOMETHING ...
    HCG_LIB_PATH=build/libhcg_x.so python bench.py ...
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hcatgnet_b200 import build as b  # noqa: E402

out = os.path.join(ROOT, sys.argv[1])
os.makedirs(os.path.dirname(out), exist_ok=True)
cmd = [b.find_nvcc(), *[f for f in b.NVCC_FLAGS if f != "-Xptxas=-v"], *sys.argv[2:], "-shared", "-o", out, *b.sources()]
subprocess.run(cmd, check=True)
print("built", out)
