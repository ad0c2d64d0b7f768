#!/usr/bin/env python
"""A/B the conv3x3 tensor-core kernel variants (conv_variant option) on batch-8 1080p: per-layer ms, interleaved."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
f = torch.from_numpy(np.stack(synth.mixed_clip(8, 1080, 1920, seed=1))).cuda()
out = torch.empty((8, 2160, 3840, 3), dtype=torch.uint8, device="cuda")
variants = sys.argv[1:] or ["1", "2"]          # "V" or "VeE": conv_variant V, epi_variant E
enh.set_option("profile", 1)
acc = {v: {} for v in variants}
for rep in range(6):
    for v in variants:
        cv, _, ev = v.partition("e")
        enh.set_option("conv_variant", int(cv)); enh.set_option("epi_variant", int(ev or 0))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); enh.upscale_device(f, mode="cnn", out=out); e1.record(); torch.cuda.synchronize()
        if rep >= 2:
            for k, ms in enh.cnn_profile_ms().items():
                acc[v].setdefault(k, []).append(ms)
            acc[v].setdefault("step", []).append(e0.elapsed_time(e1))
for v in variants:
    print("variant", v, {k: round(float(np.mean(x)), 3) for k, x in acc[v].items()})
