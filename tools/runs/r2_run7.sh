set -x
mkdir -p gpurun_out
timeout 1500 python tools/diag_conv1.py > gpurun_out/r2_run7_diag.log 2>&1; cat gpurun_out/r2_run7_diag.log
