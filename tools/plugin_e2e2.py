#!/usr/bin/env python
"""Plugin call timing variants (GPU box): result dropped vs kept alive across calls (the reference's loop keeps the previous
list alive until the next one is bound), wall clock per call, and where the time goes inside one call."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth, _lib
H, W = 1080, 1920
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
frames = [f.copy() for f in synth.mixed_clip(8, H, W, seed=4321)]
for _ in range(3): enh._process_frame_batch_gpu(frames)
def run(keep, n=20):
    out = None
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        if keep: out = enh._process_frame_batch_gpu(frames)
        else: enh._process_frame_batch_gpu(frames)
        ts.append((time.perf_counter() - t0) * 1e3)
    return ts
for keep in (False, True, False, True):
    ts = run(keep)
    print("keep=%s  mean %.2f ms  min %.2f  max %.2f  first5 %s" % (keep, np.mean(ts), min(ts), max(ts), [round(t, 1) for t in ts[:5]]))
# inside one call
t0 = time.perf_counter(); o = torch.empty((8, 2 * H, 2 * W, 3), dtype=torch.uint8, pin_memory=True); t1 = time.perf_counter()
print("fresh pinned alloc while another is alive: %.2f ms" % ((t1 - t0) * 1e3))
o2 = torch.empty((8, 2 * H, 2 * W, 3), dtype=torch.uint8, pin_memory=True); t2 = time.perf_counter()
print("second fresh pinned alloc: %.2f ms" % ((t2 - t1) * 1e3))
del o, o2
t0 = time.perf_counter(); o = torch.empty((8, 2 * H, 2 * W, 3), dtype=torch.uint8, pin_memory=True); print("after free: %.3f ms" % ((time.perf_counter() - t0) * 1e3))
