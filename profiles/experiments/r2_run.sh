timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --mode train --batch 4096 --steps 30 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/train.json
python - <<'PY'
import json
r=json.loads(open("gpurun_out/train.json").read())
print(r["value"], r["ms_per_step"], r["e2e"]["value"], {n:round(v["ms_per_launch"]*1e3,1) for n,v in r["kernels"].items()})
PY
