set -x
mkdir -p gpurun_out
./tools/ubench/mma_ts > gpurun_out/r2_mma_ts.txt 2>&1; cat gpurun_out/r2_mma_ts.txt
timeout 900 python -m pytest tests/test_rrdb_gpu.py tests/test_video_gpu.py -x -q 2>&1 | tail -5
timeout 300 python tools/rrdb_dbg.py > gpurun_out/r2_run9_rrdb_dbg.log 2>&1; grep -a rconv_dbg gpurun_out/r2_run9_rrdb_dbg.log | head -8 | cut -c1-420
timeout 900 python tools/rrdb_bench.py > gpurun_out/r2_run9_rrdb_bench.json 2> gpurun_out/r2_run9_rrdb_bench.err; cat gpurun_out/r2_run9_rrdb_bench.json; tail -3 gpurun_out/r2_run9_rrdb_bench.err
