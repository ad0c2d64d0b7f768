set -x
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -15
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
