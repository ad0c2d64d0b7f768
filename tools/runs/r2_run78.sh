timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1200 python -m pytest tests/test_cnn_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
timeout 600 python bench.py --no-cpu-baseline --no-secondary --steps 20 2>/dev/null | tail -1 | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('BENCH', l['value'], l['ms_per_step'], l['roofline']['layer_ms']['upconv1'])"
