timeout 600 python tools/enc_threads.py 2>&1 | tail -8
