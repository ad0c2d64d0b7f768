export VEX_DEBUG_PROBES=1
timeout 600 python tools/rconv_probe.py 2>&1 | tail -3
