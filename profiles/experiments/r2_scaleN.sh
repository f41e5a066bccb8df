timeout 600 python -m pytest tests/test_gpu_dp.py -x -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --mode train --steps 30 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/train2.json
python - <<'PY'
import json
r=json.loads(open("gpurun_out/train2.json").read())
print(r["n_gpus"], r["value"], r["ms_per_step"], r.get("allreduce_us"), r["e2e"]["value"])
PY
