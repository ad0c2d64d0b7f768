export VEX_DEBUG_PROBES=1
timeout 600 python tools/rconv_probe2.py 2>&1 | grep -a "PROBE\|cin=64 cout=32\|4320x7680\] " | awk '/PROBE/{a=0;b=0;print} /cout=32/{if(a++<1)print} /4320x7680/{if(b++<1)print}' | cut -c1-300
