timeout 1200 python -m pytest tests/test_cnn_gpu.py -x -q 2>&1 | tail -4
