set -x
timeout 300 python tools/conv_dbg.py 8 fuse_tail=3 2>&1 | grep -a "uptail_dbg"
