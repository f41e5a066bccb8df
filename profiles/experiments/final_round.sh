# end-of-session measurement pass: full GPU test-suite, the four bench lines, ncu launch list of the stress step
python -m pytest tests -q -m gpu 2>&1 | tail -4
python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err
python bench.py --mode train > gpurun_out/f_bench_train.json 2>> gpurun_out/f_bench.err
python bench.py --hidden 256 --convs 4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/f_stress.json 2>> gpurun_out/f_bench.err
python bench.py --hidden 256 --convs 4 --mode train --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/f_stress_train.json 2>> gpurun_out/f_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f_ref.json 2>> gpurun_out/f_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f_launches_stress.csv python bench.py --hidden 256 --convs 4 --steps 2 --warmup 1 --no-cpu-baseline --no-profile > gpurun_out/f_ncu.log 2>&1
python - <<PY
import json
for f in ("f_bench","f_bench_train","f_stress","f_stress_train","f_ref"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["value"]), d.get("ms_per_step"), "e2e", round(d["e2e"]["value"]), "packed", d.get("e2e_packed") and round(d["e2e_packed"]["value"]), d.get("roofline") and (d["roofline"]["kernel"], round(d["roofline"]["frac"],4)))
        if d.get("kernels"): print("   ", {k:(round(v["ms_per_launch"],4),v["launches_per_step"]) for k,v in list(d["kernels"].items())[:8]})
    except Exception as e: print(f, "ERR", e)
PY
tail -3 gpurun_out/f_bench.err
