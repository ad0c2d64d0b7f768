#!/usr/bin/env python
"""Which layer of the widened + deepened 2160p plan fails (GPU box)?  Runs the debug taps one by one with a sync."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import upscaler_ref as O
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2160, 3840)
model = O.make_model(0, widths=(256, 256, 512, 256), extra=2)
enh = B200VideoEnhancer(device="cuda:0", state_dict=O.export_state(model)); enh.load_models()
for kv in sys.argv[3:]:
    k, v = kv.split("="); enh.set_option(k, int(v))
f = torch.from_numpy(synth.edges_frame(H, W, np.random.default_rng(4))).cuda()[None]
for tap in (1, 2, 3, 7, 8, 4, 5, 0):
    try:
        out = enh.cnn_debug_f32(f, tap=tap); torch.cuda.synchronize()
        print("tap", tap, "ok", tuple(out.shape), float(out.abs().max()), flush=True)
        del out
    except Exception as e:
        print("tap", tap, "FAILED", str(e)[:200], flush=True)
        break
