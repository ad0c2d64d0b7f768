#!/usr/bin/env python
"""Real-ESRGAN RRDBNet (23 blocks, x4, seeded random weights) on one B200: a 532 x 532 tile (the reference's padded tile
size) in one pass, and a 1080p frame through the device tile loop (tile 512, pad 10 = 12 tiles -> 4320p).  Prints one JSON
line: ms, frames/s, algorithmic TFLOP/s against the measured bf16/fp16 dense peak, launches."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import realesrgan_ref as R
from video_enhancer_proto_b200 import B200RealESRGAN, synth

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 23
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {"bf16_tflops": 1646.9}
hd = B200RealESRGAN(device="cuda:0", scale=4, seed=0, num_block=nb)
hd._ensure_model_loaded()
for kv in sys.argv[2:]:
    k, v = kv.split("=")
    print("option", k, v, "was", hd._lib.vex_set_option(hd._ctx, k.encode(), int(v)), file=sys.stderr)
flop_px = R.flops_per_input_pixel(4, nb)
rng = np.random.default_rng(0)
res = {"num_block": nb, "flop_per_input_pixel": flop_px}

def timed(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = hd.kernel_launches()
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, (hd.kernel_launches() - l0) // reps

tile = torch.from_numpy(synth.random_frame(532, 532, rng)).cuda()[None]
ms, launches = timed(lambda: hd.upscale_device(tile), 5)
res["tile_532"] = {"ms": ms, "launches": launches, "tflops": flop_px * 532 * 532 / ms / 1e9,
                   "frac_of_burst_peak": flop_px * 532 * 532 / ms / 1e9 / peaks["bf16_tflops"]}
frame = torch.from_numpy(synth.random_frame(1080, 1920, rng)).cuda()
ms, launches = timed(lambda: hd.frame_device(frame, in_bgr=True, out_bgr=True), 3)
px = sum((py1 - py0) * (px1 - px0) for (py0, py1, px0, px1, *_r) in R.tile_plan(1080, 1920, 512, 10))
res["frame_1080p_tiled"] = {"ms": ms, "fps": 1e3 / ms, "launches": launches, "tiles": 12, "pixels_incl_padding": px,
                            "tflops": flop_px * px / ms / 1e9, "frac_of_burst_peak": flop_px * px / ms / 1e9 / peaks["bf16_tflops"]}
ms, launches = timed(lambda: hd.upscale_device(frame[None]), 2)
res["frame_1080p_whole"] = {"ms": ms, "fps": 1e3 / ms, "launches": launches, "tflops": flop_px * 1080 * 1920 / ms / 1e9}
print(json.dumps(res))
