export PYTHONPATH=.
HCG_LIB_PATH=build/libhcg_prev.so python profiles/experiments/coo_ab.py 2>&1 | tail -1
python profiles/experiments/coo_ab.py 2>&1 | tail -1
HCG_LIB_PATH=build/libhcg_prev.so python profiles/experiments/coo_ab.py 2>&1 | tail -1
python profiles/experiments/coo_ab.py 2>&1 | tail -1
