set -x
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q -k "full_size" 2>&1 | tail -5
timeout 300 python bench.py --impl reference --workload rrdb --steps 1 2>/dev/null | head -c 700
