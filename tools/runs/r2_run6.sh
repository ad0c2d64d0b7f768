set -x
mkdir -p gpurun_out
timeout 600 python tools/diag_wide.py 2160 3840 > gpurun_out/r2_run6_diag.log 2>&1; cat gpurun_out/r2_run6_diag.log | tail -12
timeout 600 python tools/diag_wide.py 2160 3840 conv1_variant=1 > gpurun_out/r2_run6_diag1.log 2>&1; cat gpurun_out/r2_run6_diag1.log | tail -12
timeout 900 python tools/rrdb_bench.py > gpurun_out/r2_run6_rrdb_bench.json 2> gpurun_out/r2_run6_rrdb_bench.err; cat gpurun_out/r2_run6_rrdb_bench.json; tail -3 gpurun_out/r2_run6_rrdb_bench.err
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q 2>&1 | tail -3
