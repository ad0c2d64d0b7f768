#!/usr/bin/env python
"""Where the time of rconv3x3_kernel goes (VEX_DEBUG_PROBES build only, wrong pixels): a whole 1080p frame through the
23-block x4 network with the activation loads, the epilogue and the MMAs switched off in turn."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200RealESRGAN, synth

hd = B200RealESRGAN(device="cuda:0", scale=4, seed=0, num_block=23)
hd._ensure_model_loaded()
frame = torch.from_numpy(synth.random_frame(1080, 1920, np.random.default_rng(0))).cuda()[None]
res = {}
for probe in (0, 256, 512, 768, 2048, 2048 + 256, 2048 + 512, 2048 + 768, 0):
    rc = hd.set_option("thin_probe", probe)
    assert rc >= 0, rc
    hd.upscale_device(frame); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(2): hd.upscale_device(frame)
    e1.record(); torch.cuda.synchronize()
    res[f"probe_{probe}" + ("_again" if f"probe_{probe}" in res else "")] = round(e0.elapsed_time(e1) / 2, 2)
print(json.dumps(res))
