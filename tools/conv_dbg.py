#!/usr/bin/env python
"""Print conv_tc per-role cycle accounting for one batch-8 1080p forward (GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
f = torch.from_numpy(np.stack(synth.mixed_clip(n, 1080, 1920, seed=1))).cuda()
for _ in range(2): enh.upscale_device(f, mode="cnn")
torch.cuda.synchronize()
for kv in sys.argv[2:]:
    k, v = kv.split("="); enh.set_option(k, int(v))
enh.set_option("conv_dbg", 1)
enh.upscale_device(f, mode="cnn"); torch.cuda.synchronize()
