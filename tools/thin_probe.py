#!/usr/bin/env python
"""Which role bounds the im2col27 kernels?  Per-layer ms with timing probes (wrong results on purpose)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
f = torch.from_numpy(np.stack(synth.mixed_clip(8, 1080, 1920, seed=1))).cuda()
out = torch.empty((8, 2160, 3840, 3), dtype=torch.uint8, device="cuda")
enh.set_option("profile", 1)
for probe in ([int(a) for a in sys.argv[1:]] or (0, 1, 2, 4, 3, 5, 6, 7, 0)):
    enh.set_option("thin_probe", probe)
    acc = []
    for rep in range(4):
        enh.upscale_device(f, mode="cnn", out=out); torch.cuda.synchronize()
        if rep: acc.append(enh.cnn_profile_ms())
    print("probe", probe, {k: round(float(np.mean([a[k] for a in acc])), 3) for k in ("conv1", "upconv1", "upconv2")}, flush=True)
