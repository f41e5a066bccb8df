set -x
timeout 900 python bench.py 2>&1 | tail -1 > gpurun_out/bench_default.json
python - <<'PY'
import json
r=json.loads(open("gpurun_out/bench_default.json").read())
print(r["value"], r["ms_per_step"], r["e2e"]["value"], r.get("e2e_packed",{}).get("value"), r["gpu_launches"], r["roofline"])
print(r["kernels"])
print(r["train"]["value"], r["stress"]["value"], r["ensemble"])
PY
