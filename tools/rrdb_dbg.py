#!/usr/bin/env python
"""Per-role cycle accounting of the RRDB conv kernel: one 1-block network on a 532 x 532 tile (GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, ctypes as C
from video_enhancer_proto_b200 import B200RealESRGAN, synth
H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (532, 532)
hd = B200RealESRGAN(device="cuda:0", scale=4, seed=0, num_block=1); hd._ensure_model_loaded()
x = torch.from_numpy(synth.random_frame(H, W, np.random.default_rng(0))).cuda()[None]
for _ in range(2): hd.upscale_device(x)
torch.cuda.synchronize()
if len(sys.argv) > 3: print("probe", sys.argv[3], hd._lib.vex_set_option(hd._ctx, b"thin_probe", int(sys.argv[3])), file=sys.stderr)
hd._lib.vex_set_option(hd._ctx, b"conv_dbg", 1)
hd.upscale_device(x); torch.cuda.synchronize()
