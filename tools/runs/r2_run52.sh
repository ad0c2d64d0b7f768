timeout 600 python tools/conv_iss_ab.py 2>&1 | grep -a "issuers\|Error\|error" | cut -c1-420
