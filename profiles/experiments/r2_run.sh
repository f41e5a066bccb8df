export PYTHONPATH=.
python profiles/experiments/stress_ab.py 2>&1 | tail -4
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
