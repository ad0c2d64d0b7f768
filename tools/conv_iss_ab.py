#!/usr/bin/env python
"""Same-run A/B of the second MMA-issuing thread of conv3x3_tc_kernel v3 (option conv_issuers: 0 = two for BN = 128, 1 = one):
sustained 8-frame 1080p steps, alternating; ms per step and the per-layer means; the outputs must be identical."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
f = torch.from_numpy(np.stack(synth.mixed_clip(8, 1080, 1920, seed=1))).cuda()
out = torch.empty((8, 2160, 3840, 3), dtype=torch.uint8, device="cuda")
ref = None
for rnd in range(4):
    for one in (0,):
        enh.set_option("conv_issuers", one)
        for _ in range(5):
            enh.upscale_device(f, mode="cnn", out=out)
        torch.cuda.synchronize()
        if ref is None: ref = out.clone()
        assert torch.equal(out, ref)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(40):
            enh.upscale_device(f, mode="cnn", out=out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 40
        enh.set_option("profile", 1)
        for _ in range(5): enh.upscale_device(f, mode="cnn", out=out)
        torch.cuda.synchronize()
        st = enh.cnn_profile_stats(); enh.set_option("profile", 0)
        mean = st["mean"] if isinstance(st, dict) and "mean" in st else st
        print(f"issuers={'one' if one else 'two'}  {ms:7.3f} ms/step  {8e3 / ms:6.1f} fps  {mean}", flush=True)
