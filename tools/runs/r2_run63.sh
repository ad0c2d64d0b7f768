timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 2>/dev/null | tail -1 > gpurun_out/r2_bench_v4.json; python -c "
import json; l=json.loads(open('gpurun_out/r2_bench_v4.json').read()); print('BENCH', l['value'], l['ms_per_step'], l['e2e']['value'], l['roofline']['frac'], l['roofline']['layer_ms'], l['clocks'], l['gpu_launches'])"
