timeout 600 python tools/tail4_check.py 2>&1 | grep -a "diff\|fuse_tail\|uptail_dbg\|Error\|error" | cut -c1-600
