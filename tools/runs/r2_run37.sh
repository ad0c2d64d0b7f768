set -x
timeout 1500 python -m pytest tests/test_rrdb_gpu.py -x -q 2>&1 | tail -12
timeout 600 python tools/rrdb_bench.py 23 2>/dev/null | tail -1
timeout 600 python tools/rrdb_bench.py 23 rrdb_zigzag=0 2>/dev/null | tail -1
timeout 600 python tools/rrdb_bench.py 23 2>/dev/null | tail -1
