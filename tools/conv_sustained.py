#!/usr/bin/env python
"""Sustained (power-capped) step time per conv kernel variant: 20 back-to-back batch-8 1080p steps each, no profiling."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
f = torch.from_numpy(np.stack(synth.mixed_clip(8, 1080, 1920, seed=1))).cuda()
out = torch.empty((8, 2160, 3840, 3), dtype=torch.uint8, device="cuda")
for v in (sys.argv[1:] or ["1", "2", "3"]):
    cv, _, ev = v.partition("e")
    enh.set_option("conv_variant", int(cv)); enh.set_option("epi_variant", int(ev or 0))
    for _ in range(3): enh.upscale_device(f, mode="cnn", out=out)
    torch.cuda.synchronize()
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(21)]
    ev0[0].record()
    for i in range(20):
        enh.upscale_device(f, mode="cnn", out=out); ev0[i + 1].record()
    torch.cuda.synchronize()
    ms = [ev0[i].elapsed_time(ev0[i + 1]) for i in range(20)]
    print(f"variant {v}: first 3 steps {np.mean(ms[:3]):.2f} ms, last 10 steps {np.mean(ms[10:]):.2f} ms -> {8e3 / np.mean(ms[10:]):.0f} fps sustained", flush=True)
    time.sleep(3)
