#!/usr/bin/env python
"""The reference's OWN GPU path timed on the same B200 (last paragraph; VERDICT r1 missing #6).

Not the product and never imported by it: this is the reference's code as written (`AdvancedUpscaler` through
torch/cuDNN, `_process_frame_batch_gpu` with its pageable fp32 H2D and per-frame fp32 D2H, app.py:529-553), run on
cuda:0 with the oracle's module (bit-identical to the reference class, oracle/make_golden.py), so that the CPU arm's
ratio is not the only comparison: it is the one "existing GPU kernel" the reference has.  Prints one JSON line.

    python tools/ref_cuda_path.py [--frames 8] [--steps 5]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import upscaler_ref as O                      # noqa: E402
from video_enhancer_proto_b200 import synth                # noqa: E402

H, W = 1080, 1920


def as_written(model, frames_batch):
    """app.py:529-553 verbatim in structure: host normalise + stack, pageable H2D, model, per-frame fp32 D2H + clip + swap."""
    import cv2
    batch_tensor = []
    for frame in frames_batch:
        frame_tensor = torch.from_numpy(frame).permute(2, 0, 1).float() / 255.0
        batch_tensor.append(frame_tensor)
    batch_tensor = torch.stack(batch_tensor).to("cuda")
    with torch.no_grad():
        enhanced_batch = model(batch_tensor)
    out = []
    for i in range(enhanced_batch.shape[0]):
        enhanced_frame = enhanced_batch[i].permute(1, 2, 0).cpu().numpy()
        enhanced_frame = np.clip(enhanced_frame * 255, 0, 255).astype(np.uint8)
        out.append(cv2.cvtColor(enhanced_frame, cv2.COLOR_RGB2BGR))
    return out


def timed(fn, steps, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    frames = synth.mixed_clip(a.frames, H, W, seed=1234)
    model = O.make_model(0).cuda().eval()
    x = torch.stack([torch.from_numpy(f).permute(2, 0, 1).float() / 255.0 for f in frames]).cuda()
    res = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__, "frames_per_step": a.frames,
           "what": "reference AdvancedUpscaler through torch/cuDNN on this GPU (tools only; not the product path)"}

    def fwd():
        with torch.no_grad():
            model(x)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    res["fp32_cudnn_forward_fps"] = a.frames / timed(fwd, a.steps)
    torch.backends.cudnn.allow_tf32 = True
    res["tf32_cudnn_forward_fps"] = a.frames / timed(fwd, a.steps)
    torch.backends.cudnn.benchmark = True
    m2 = O.make_model(0).cuda().eval().to(memory_format=torch.channels_last)
    x2 = x.to(memory_format=torch.channels_last)

    def fwd_bf16():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            m2(x2)
    res["bf16_autocast_channels_last_forward_fps"] = a.frames / timed(fwd_bf16, a.steps)
    torch.backends.cudnn.allow_tf32 = False
    res["as_written_end_to_end_fps"] = a.frames / timed(lambda: as_written(model, frames), max(2, a.steps // 2), warmup=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
