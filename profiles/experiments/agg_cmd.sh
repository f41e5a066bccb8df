python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "aggregation or backward_paths or tensor_core_gemm" 2>&1 | tail -2
python bench.py --hidden 256 --convs 4 --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('stress', round(d['value']), d['ms_per_step']); print({k:(round(v['ms_per_launch'],4),v['launches_per_step']) for k,v in list(d['kernels'].items())[:4]}); print({k:(round(v['achieved']),round(v['frac'],3)) for k,v in d['aggregation_roofline'].items()})"
python profiles/experiments/bench_cfg.py 0=1 -- --no-cpu-baseline --steps 20 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('layer-wise H=64', round(d['value']), d['ms_per_step']); print({k:(round(v['ms_per_launch'],4),v['launches_per_step']) for k,v in list(d['kernels'].items())[:4]}); print(d['aggregation_roofline'])"
