set -x
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q 2>&1 | tail -8
timeout 600 python tools/rrdb_bench.py 23 2>/dev/null | tail -1 | cut -c1-700
timeout 300 python tools/rrdb_dbg.py 1080 1920 2>&1 | grep -a "rconv_dbg" | sed -n '1,5p;16,20p' | cut -c1-330
