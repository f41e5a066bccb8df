set -x
PYTHONPATH=. timeout 600 python profiles/experiments/ensemble_rate.py 2>&1 | tail -12
timeout 600 python -m pytest tests/test_gpu_ensemble.py tests/test_gpu_module.py -x -q 2>&1 | tail -5
