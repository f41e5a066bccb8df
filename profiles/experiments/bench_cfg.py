"""Run bench.py with library switches preset: python profiles/experiments/bench_cfg.py KEY=VALUE ... -- <bench args>
(KEY is the integer HCG_CFG_* index).  Used for A/B timing of kernel paths; not part of the product."""
import runpy
import sys

sys.path.insert(0, ".")
from hcatgnet_b200 import _lib

args = sys.argv[1:]
split = args.index("--")
for kv in args[:split]:
    k, v = kv.split("=")
    _lib.config_set(int(k), int(v))
sys.argv = ["bench.py"] + args[split + 1:]
runpy.run_path("bench.py", run_name="__main__")
