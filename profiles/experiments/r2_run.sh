set -x
for T in 2 3; do
  HCG_TC_SLOTS=$T timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tc.py -x -q -m gpu -k "forward or tc or cross" 2>&1 | tail -3
  HCG_TC_SLOTS=$T timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('SLOTS=$T', d['value'], d['ms_per_step'], {k:round(v['ms_per_launch']*1e3,1) for k,v in d['kernels'].items()})"
done
HCG_TC_SLOTS=3 HCG_LIB_PATH=build/libhcg_timing.so timeout 200 python profiles/tc_phase_timing.py 2>&1
