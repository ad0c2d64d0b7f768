set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cnn_gpu.py -x -q -k "fused_tail" > gpurun_out/r2_run15_tail.log 2>&1; tail -6 gpurun_out/r2_run15_tail.log
timeout 300 python tools/conv_dbg.py 8 fuse_tail=3 > gpurun_out/r2_run15_dbg3.log 2>&1; grep -a "uptail_dbg" gpurun_out/r2_run15_dbg3.log
timeout 300 python tools/conv_dbg.py 8 > gpurun_out/r2_run15_dbg1.log 2>&1; grep -a "uptail_dbg" gpurun_out/r2_run15_dbg1.log
