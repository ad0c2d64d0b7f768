export VEX_DEBUG_PROBES=1
for pr in 0 2816; do
timeout 300 python tools/rrdb_dbg.py 1080 1920 $pr 2>&1 | grep -a "rconv_dbg\|probe" | head -20 | cut -c1-330
done
