set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "backward" 2>&1 | tail -4
timeout 600 python bench.py --mode train --steps 30 --warmup 5 --no-cpu-baseline 2>gpurun_out/tb.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('train', d['value'], d['ms_per_step'], {k:round(v['ms_per_launch']*1e3,1) for k,v in d['kernels'].items()})"; tail -3 gpurun_out/tb.err
HCG_LIB_PATH=build/libhcg_timing.so python profiles/tb_phase_timing.py 2>&1 | tail -17
