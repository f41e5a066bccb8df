python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 2>&1 | tail -1 > gpurun_out/scale8.json
python - <<'PY'
import json
r=json.loads(open("gpurun_out/scale8.json").read())
print(r["n_gpus"], r["value"], r["ms_per_step"], r["e2e"]["value"], r["e2e_packed"]["value"])
print(r["train"]["value"], r["train"].get("allreduce_us"), r["stress"]["value"], r["ensemble"]["value"])
PY
