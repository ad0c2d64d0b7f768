#!/usr/bin/env python
"""End-to-end fps of vex_upscale_host (pinned host in/out) against frames per call and frames per chunk."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth, _lib
H, W = 1080, 1920
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
lib = _lib.load()
pool = synth.mixed_clip(12, H, W, seed=1234)
fin = H * W * 3
for frames, chunk in [(48, 8), (48, 4), (48, 16), (48, 6), (96, 8), (48, 8)]:
    pin_in = enh._pinned_buf("in", frames * fin); pin_out = enh._pinned_buf("out", frames * fin * 4)
    vin = pin_in.numpy()[: frames * fin].reshape(frames, H, W, 3)
    for i in range(frames): vin[i] = pool[i % 12]
    def step():
        _lib.check(enh._ctx, lib.vex_upscale_host(enh._ctx, _lib.VEX_MODE_CNN, pin_in.data_ptr(), pin_out.data_ptr(), frames, H, W, chunk, 0), "host")
    step(); step()
    t0 = time.perf_counter()
    for _ in range(6): step()
    dt = time.perf_counter() - t0
    print(f"frames/call {frames:3d} chunk {chunk:2d}: {6 * frames / dt:7.1f} fps  ({dt / 6 * 1e3:.1f} ms per call)", flush=True)
