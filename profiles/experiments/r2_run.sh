python -c "import __graft_entry__ as g; g.build(); g.smoke()" 2>&1 | tail -1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 900 python bench.py 2>&1 | tail -1 > gpurun_out/r2_final_bench.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 > gpurun_out/r2_final_reference.json
python - <<'PY'
import json
r=json.loads(open("gpurun_out/r2_final_bench.json").read())
print(r["value"], r["ms_per_step"], r["e2e"]["value"], r["e2e_packed"]["value"], r["train"]["value"], r["stress"]["value"], r["ensemble"]["value"])
print(json.loads(open("gpurun_out/r2_final_reference.json").read())["value"])
PY
