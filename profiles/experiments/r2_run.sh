set -x
for k in 1 2; do
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sub 2>&1 | tail -1 > gpurun_out/b$k.json
done
HCG_BENCH_CSR=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sub 2>&1 | tail -1 > gpurun_out/b3.json
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sub --no-profile 2>&1 | tail -1 > gpurun_out/b4.json
python - <<'PY'
import json
for k in (1,2,3,4):
    r=json.loads(open(f"gpurun_out/b{k}.json").read())
    print(k, r["value"], r["e2e"]["value"], r["e2e"].get("wall_ms_per_step"), r.get("e2e_packed",{}).get("value"), r.get("e2e_packed",{}).get("ms_per_step"))
PY
