timeout 900 python bench.py 2>&1 | tail -1 > gpurun_out/r2_final_bench.json
python - <<'PY'
import json
r=json.loads(open("gpurun_out/r2_final_bench.json").read())
print(r["value"], r["ms_per_step"], r["e2e"]["value"], r["e2e_packed"]["value"], r["train"]["value"], r["train"]["e2e"]["value"], r["stress"]["value"], r["ensemble"]["value"])
print(r["train"]["kernels"])
PY
