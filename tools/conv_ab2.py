#!/usr/bin/env python
"""Sustained A/B of option sets on the batch-8 1080p step (20 back-to-back steps each, interleaved twice): name=value pairs
joined by commas per variant, e.g.  python tools/conv_ab2.py default epi_variant=8 conv3_bn=128"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
f = torch.from_numpy(np.stack(synth.mixed_clip(8, 1080, 1920, seed=1))).cuda()
out = torch.empty((8, 2160, 3840, 3), dtype=torch.uint8, device="cuda")
variants = sys.argv[1:] or ["default"]
for rnd in range(2):
    for v in variants:
        opts = [] if v == "default" else [kv.split("=") for kv in v.split(",")]
        prev = [(k, enh.set_option(k, int(x))) for k, x in opts]
        for _ in range(3): enh.upscale_device(f, mode="cnn", out=out)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(21)]
        ev[0].record()
        for i in range(20):
            enh.upscale_device(f, mode="cnn", out=out); ev[i + 1].record()
        torch.cuda.synchronize()
        ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(20)]
        print(f"round {rnd} {v}: last 10 steps {np.mean(ms[10:]):.2f} ms -> {8e3 / np.mean(ms[10:]):.0f} fps", flush=True)
        for k, p in prev: enh.set_option(k, p)
        time.sleep(2)
