#!/usr/bin/env python
"""Epilogue cost breakdown of rconv3x3_kernel (VEX_DEBUG_PROBES build, wrong pixels): conv1-type layers and the heads at
1080p with the shuffles (4096) / the global stores (8192) switched off."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200RealESRGAN, synth
hd = B200RealESRGAN(device="cuda:0", scale=4, seed=0, num_block=1); hd._ensure_model_loaded()
x = torch.from_numpy(synth.random_frame(1080, 1920, np.random.default_rng(0))).cuda()[None]
for _ in range(2): hd.upscale_device(x)
torch.cuda.synchronize()
hd.set_option("conv_dbg", 1)
for probe in (0, 4096, 8192, 12288):
    print("PROBE", probe, hd.set_option("thin_probe", probe), file=sys.stderr, flush=True)
    hd.upscale_device(x); torch.cuda.synchronize()
