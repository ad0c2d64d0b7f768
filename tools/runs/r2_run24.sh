timeout 600 python tools/conv_ab2.py default epi_variant=8 conv3_bn=128 2>&1 | grep "round"
