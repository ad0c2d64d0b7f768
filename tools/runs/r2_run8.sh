set -x
mkdir -p gpurun_out
timeout 1500 python tools/diag_conv1.py > gpurun_out/r2_run8_diag.log 2>&1; cat gpurun_out/r2_run8_diag.log | head -12
timeout 300 python tools/rrdb_dbg.py > gpurun_out/r2_run8_rrdb_dbg.log 2>&1; grep -a rconv_dbg gpurun_out/r2_run8_rrdb_dbg.log
timeout 1500 python -m pytest tests -m gpu -x -q -k "not all_300" > gpurun_out/r2_run8_pytest.log 2>&1; tail -8 gpurun_out/r2_run8_pytest.log
