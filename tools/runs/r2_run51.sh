set -x
timeout 1200 python -m pytest tests/test_cnn_gpu.py tests/test_async_gpu.py -x -q 2>&1 | tail -4
timeout 900 python bench.py --no-cpu-baseline --no-secondary --steps 20 2>/dev/null | tail -1 | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('BENCH', l['value'], l['ms_per_step'], l['roofline']['frac'], l['roofline']['layer_ms'], l['clocks'])"
