N=${HCG_N:-2}
timeout 600 python -m pytest tests/test_gpu_dp.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 2>&1 | tail -1 > gpurun_out/scale$N.json
python - <<'PY'
import json, os
n=os.environ.get("HCG_N","2")
r=json.loads(open(f"gpurun_out/scale{n}.json").read())
print(r["n_gpus"], r["value"], r["ms_per_step"], r["e2e"]["value"], r["e2e_packed"]["value"])
print(r["train"]["value"], r["train"].get("allreduce_us"), r["stress"]["value"], r["ensemble"]["value"])
PY
