# round 2, second GPU call: RRDB parity, uptail per-role cycles, plugin path breakdown, remaining new tests
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q > gpurun_out/r2_run2_rrdb.log 2>&1; tail -25 gpurun_out/r2_run2_rrdb.log
timeout 300 python tools/conv_dbg.py 8 > gpurun_out/r2_run2_dbg.log 2>&1; grep -a "dbg" gpurun_out/r2_run2_dbg.log
timeout 300 python tools/plugin_e2e.py > gpurun_out/r2_run2_plugin.log 2>&1; cat gpurun_out/r2_run2_plugin.log
timeout 1500 python -m pytest tests -m gpu -x -q -k "not all_300" --deselect tests/test_rrdb_gpu.py > gpurun_out/r2_run2_pytest.log 2>&1; tail -15 gpurun_out/r2_run2_pytest.log
