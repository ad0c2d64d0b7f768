set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_filters_gpu.py tests/test_fuzz_gpu.py -x -q 2>&1 | tail -4
timeout 600 python tools/filters_bench.py > gpurun_out/r2_filters.json 2> gpurun_out/r2_filters.err; tail -2 gpurun_out/r2_filters.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_filters.json'))
for k,v in d['stages'].items(): print(k, round(v['fps']), round(v.get('hbm_frac',0),3), v.get('max_diff'), v.get('mismatch_frac'))
PY
timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-file > gpurun_out/r2_run17_bench.json 2> gpurun_out/r2_run17_bench.err; tail -2 gpurun_out/r2_run17_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run17_bench.json'))
print(d['value'], d['e2e']['value'], d['e2e']['streaming']['value'], d['e2e_plugin']['value'], d['bicubic']['e2e'])
PY
