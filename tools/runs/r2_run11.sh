set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cnn_gpu.py -x -q -k "fused_tail or taps_and_output or golden" > gpurun_out/r2_run11_tail.log 2>&1; tail -12 gpurun_out/r2_run11_tail.log
timeout 300 python tools/conv_dbg.py 8 > gpurun_out/r2_run11_dbg.log 2>&1; grep -a "uptail_dbg" gpurun_out/r2_run11_dbg.log
timeout 300 python tools/conv_dbg.py 8 fuse_tail=2 > gpurun_out/r2_run11_dbg2.log 2>&1; grep -a "uptail_dbg" gpurun_out/r2_run11_dbg2.log
timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-file --no-secondary > gpurun_out/r2_run11_bench.json 2> gpurun_out/r2_run11_bench.err; tail -3 gpurun_out/r2_run11_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run11_bench.json'))
print(d['value'], d['e2e']['value'], d['roofline']['layer_ms'], d['roofline']['frac'], d['roofline']['whole_net_frac'], d['clocks'])
PY
timeout 900 python tools/rrdb_bench.py 23 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('tile %.2f ms' % d['tile_532']['ms'], 'tiled %.1f ms' % d['frame_1080p_tiled']['ms'], 'whole %.1f ms' % d['frame_1080p_whole']['ms'])"
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q 2>&1 | tail -3
