# end-of-round measurement pass (third part of round 2): full GPU test-suite, default bench line, reference arm, launch list of the training step
export PYTHONPATH=$PWD
mkdir -p gpurun_out
python -m pytest tests -q -m gpu 2>&1 | tail -3
python bench.py > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/c_ref.json 2>> gpurun_out/c_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c_launches_train.csv python bench.py --mode train --steps 2 --warmup 1 --no-cpu-baseline --no-profile --no-sub > gpurun_out/c_ncu.log 2>&1
python - <<PY
import json
for f in ("c_bench","c_ref"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["value"]), d.get("ms_per_step"), "e2e", round(d["e2e"]["value"]), "packed", d.get("e2e_packed") and round(d["e2e_packed"]["value"]), d.get("roofline") and (d["roofline"]["kernel"], round(d["roofline"]["frac"],4)))
        for sub in ("train","stress","ensemble"):
            if d.get(sub): print("  ", sub, round(d[sub]["value"]), d[sub].get("ms_per_step"))
        if d.get("kernels"): print("   ", {k:(round(v["ms_per_launch"],4),v["launches_per_step"]) for k,v in list(d["kernels"].items())[:8]})
    except Exception as e: print(f, "ERR", e)
PY
tail -3 gpurun_out/c_bench.err
