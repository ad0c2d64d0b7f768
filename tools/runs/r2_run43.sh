set -x
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q 2>&1 | tail -8
timeout 600 python tools/rrdb_bench.py 23 2>/dev/null | tail -1 | cut -c1-700
timeout 600 python tools/rrdb_bench.py 23 rrdb_half=0 2>/dev/null | tail -1 | cut -c1-700
