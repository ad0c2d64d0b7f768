set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "not all_300" > gpurun_out/r2_run4_pytest.log 2>&1; tail -8 gpurun_out/r2_run4_pytest.log
timeout 300 python tools/conv_dbg.py 8 > gpurun_out/r2_run4_dbg.log 2>&1; grep -a "_dbg" gpurun_out/r2_run4_dbg.log
timeout 300 python tools/conv_dbg.py 8 conv1_variant=1 > gpurun_out/r2_run4_dbg1.log 2>&1; grep -a "im2col27_dbg" gpurun_out/r2_run4_dbg1.log
timeout 900 python tools/rrdb_bench.py > gpurun_out/r2_run4_rrdb_bench.json 2> gpurun_out/r2_run4_rrdb_bench.err; cat gpurun_out/r2_run4_rrdb_bench.json; tail -3 gpurun_out/r2_run4_rrdb_bench.err
