export PYTHONPATH=.
python profiles/experiments/stress_ab.py 2>&1 | tail -1
for m in 5 6 8; do HCG_LIB_PATH=build/libhcg_m$m.so python profiles/experiments/stress_ab.py 2>&1 | tail -1; done
