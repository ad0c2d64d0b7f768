set -x
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q -s -k "full_size" 2>&1 | grep -v "^$" | tail -15
