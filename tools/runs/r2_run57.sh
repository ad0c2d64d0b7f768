timeout 1200 python -m pytest tests/test_cnn_gpu.py tests/test_async_gpu.py tests/test_video_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -4
timeout 900 python bench.py --no-cpu-baseline --steps 20 2>/dev/null | tail -1 > gpurun_out/r2_bench_v3.json; python -c "
import json; l=json.loads(open('gpurun_out/r2_bench_v3.json').read()); print('BENCH', l['value'], l['ms_per_step'], l['e2e']['value'], l['roofline']['frac'], l['roofline']['layer_ms'], l['clocks'])"
