timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for k in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-sub 2>&1 | tail -1 > gpurun_out/b$k.json; done
python - <<'PY'
import json
for k in (1,2):
    r=json.loads(open(f"gpurun_out/b{k}.json").read())
    print(r["value"], r["ms_per_step"], {n:round(v["ms_per_launch"]*1e3,1) for n,v in r["kernels"].items()}, r["e2e_packed"]["value"], r["gpu_launches"])
PY
