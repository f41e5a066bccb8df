set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sub 2>&1 | tail -1 > gpurun_out/b1.json
python - <<'PY'
import json
r=json.loads(open("gpurun_out/b1.json").read())
print(r["value"], r["e2e"]["value"], r["e2e_packed"])
PY
