set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_final_pytest.log 2>&1; tail -3 gpurun_out/r2_final_pytest.log
python __graft_entry__.py --smoke > gpurun_out/r2_final_smoke.log 2>&1; tail -2 gpurun_out/r2_final_smoke.log
( time python bench.py --steps 20 > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err ) 2>&1 | grep real
( time python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r2_final_bench_reference.json 2>> gpurun_out/r2_final_bench.err ) 2>&1 | grep real
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_final_bench.json'))
r=d['roofline']
print('value', d['value'], 'e2e', d['e2e']['value'], 'stream', d['e2e']['streaming']['value'], 'plugin', d['e2e_plugin']['value'], 'file', d['e2e_file']['value'])
print('frac', r['frac'], 'whole', r['whole_net_frac'], 'thin', r['thin']['ms_per_step'], {k: round(v,3) for k,v in r['layer_ms'].items()}, d['clocks'])
PY
