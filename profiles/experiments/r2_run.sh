set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -6
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
