#!/usr/bin/env python
"""tail_tc3.cuh (option fuse_tail = 4) against the default fused tail: output difference on full-range weights, then an
alternating A/B of ms per 8-frame 1080p step and of the tail's own time."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import upscaler_ref as O
from video_enhancer_proto_b200 import B200VideoEnhancer, synth

enh = B200VideoEnhancer(device="cuda:0", state_dict=O.export_state(O.make_stress_model())); enh.load_models()
for (h, w, n) in ((67, 94, 2), (360, 480, 3), (1080, 1920, 2)):
    f = torch.from_numpy(np.stack(synth.mixed_clip(n, h, w, seed=3))).cuda()
    enh.set_option("fuse_tail", 1); a = enh.upscale_device(f, out_bgr=True).cpu().numpy().astype(int)
    enh.set_option("fuse_tail", 4); b = enh.upscale_device(f, out_bgr=True).cpu().numpy().astype(int)
    d = np.abs(a - b)
    print(f"{h}x{w}: max |diff| {d.max()}  differing {float((d > 0).mean()):.2e}  range {a.min()}..{a.max()}", flush=True)
f = torch.from_numpy(np.stack(synth.mixed_clip(8, 1080, 1920, seed=1))).cuda()
out = torch.empty((8, 2160, 3840, 3), dtype=torch.uint8, device="cuda")
for rnd in range(3):
    for fuse in (1, 4):
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
        print(f"fuse_tail={fuse}  {ms:7.3f} ms/step  {8e3 / ms:6.1f} fps   tail {pr['upconv1'] + pr['upconv2']:.2f} ms", flush=True)
enh.set_option("conv_dbg", 1); enh.set_option("fuse_tail", 4)
enh.upscale_device(f, mode="cnn", out=out); torch.cuda.synchronize()
