set -x
timeout 900 python -m pytest tests/test_filters_gpu.py tests/test_fuzz_gpu.py -x -q 2>&1 | tail -4
timeout 600 python tools/filters_bench.py > gpurun_out/r2_filters2.json 2> gpurun_out/r2_filters2.err; tail -2 gpurun_out/r2_filters2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_filters2.json'))
for k,v in d['stages'].items(): print(k, round(v['fps']), round(v.get('hbm_frac',0),3), v.get('max_diff'), v.get('mismatch_frac'))
PY
