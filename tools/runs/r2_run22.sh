timeout 300 python tools/rrdb_dbg.py 2>&1 | grep -a rconv_dbg | head -7 | cut -c1-420
