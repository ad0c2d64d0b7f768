set -x
timeout 1500 python -m pytest tests/test_rrdb_gpu.py -x -q 2>&1 | tail -12
timeout 600 python tools/rrdb_bench.py 23 2>/dev/null | tail -1
timeout 600 python tools/rrdb_bench.py 23 rrdb_stages=2 2>/dev/null | tail -1
timeout 300 python tools/rrdb_dbg.py 1080 1920 2>&1 | grep -a "rconv_dbg" | sed -n '1,6p;19,26p' | cut -c1-330
