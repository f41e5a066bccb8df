timeout 600 python -m pytest tests/test_gpu_edge_list.py -x -q 2>&1 | tail -3
PYTHONPATH=. timeout 300 python profiles/experiments/packed_fwd_time.py 2>&1 | tail -3
