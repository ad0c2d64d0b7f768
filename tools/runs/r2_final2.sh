set -x
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 2>/dev/null | tail -1 > gpurun_out/r2_bench_final.json
timeout 900 python bench.py --workload rrdb 2>/dev/null | tail -1 > gpurun_out/r2_bench_rrdb_final.json
timeout 600 python bench.py --impl reference --steps 1 2>/dev/null | tail -1 > gpurun_out/r2_bench_reference_final.json
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_final.json','gpurun_out/r2_bench_rrdb_final.json','gpurun_out/r2_bench_reference_final.json'):
    l=json.loads(open(f).read()); print(f, l['value'], l.get('ms_per_step'), l['e2e']['value'], l.get('roofline',{}).get('frac'), l.get('clocks'))
PY
