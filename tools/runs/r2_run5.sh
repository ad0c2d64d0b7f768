set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "not all_300" > gpurun_out/r2_run5_pytest.log 2>&1; tail -8 gpurun_out/r2_run5_pytest.log
timeout 300 python tools/conv_dbg.py 8 > gpurun_out/r2_run5_dbg.log 2>&1; grep -a "im2col27_dbg\|uptail_dbg" gpurun_out/r2_run5_dbg.log
timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-file > gpurun_out/r2_run5_bench.json 2> gpurun_out/r2_run5_bench.err; tail -3 gpurun_out/r2_run5_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run5_bench.json'))
print(d['value'], d['e2e']['value'], d['roofline']['layer_ms'], d['roofline']['frac'], d['roofline']['whole_net_frac'], d['e2e_plugin']['value'], d['clocks'])
PY
bash tools/sanitize.sh r2 2>&1 | grep -v "^+" | tail -12
