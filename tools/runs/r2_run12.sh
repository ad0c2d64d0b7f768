set -x
timeout 300 python tools/plugin_e2e2.py 2>&1 | tail -12
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_run12_pytest.log 2>&1; tail -4 gpurun_out/r2_run12_pytest.log
