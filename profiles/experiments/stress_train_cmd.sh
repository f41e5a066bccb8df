python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -4
python bench.py --hidden 256 --convs 4 --mode train --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/stress_tr.json 2> gpurun_out/stress_tr.err
python - <<PY
import json
d=json.loads(open("gpurun_out/stress_tr.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"]); print({k:(round(v["ms_per_launch"],4),v["launches_per_step"]) for k,v in d["kernels"].items()})
PY
tail -3 gpurun_out/stress_tr.err
