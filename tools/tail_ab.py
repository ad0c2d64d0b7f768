#!/usr/bin/env python
"""A/B of the fused tail (option fuse_tail): 8-frame 1080p steps, sustained, alternating; ms per step and tail ms."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
f = torch.from_numpy(np.stack(synth.mixed_clip(8, 1080, 1920, seed=1))).cuda()
out = torch.empty((8, 2160, 3840, 3), dtype=torch.uint8, device="cuda")
for rnd in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    for fuse in (1, 0):
        enh.set_option("fuse_tail", fuse)
        for _ in range(5):
            enh.upscale_device(f, mode="cnn", out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30):
            enh.upscale_device(f, mode="cnn", out=out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 30
        enh.set_option("profile", 1)
        enh.upscale_device(f, mode="cnn", out=out); torch.cuda.synchronize()
        pr = enh.cnn_profile_ms(); enh.set_option("profile", 0)
        print(f"fuse_tail={fuse}  {ms:7.3f} ms/step  {8e3 / ms:6.1f} fps   tail {pr['upconv1'] + pr['upconv2']:.2f} ms  "
              + " ".join(f"{k}={v:.2f}" for k, v in pr.items()), flush=True)
