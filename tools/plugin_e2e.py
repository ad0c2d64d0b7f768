#!/usr/bin/env python
"""Where the time of the plugin call `_process_frame_batch_gpu` goes (GPU box): the method, the C call alone with a
preallocated pinned output, the pinned allocation alone, and the device-resident step for scale."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from video_enhancer_proto_b200 import B200VideoEnhancer, synth, _lib
H, W = 1080, 1920
enh = B200VideoEnhancer(device="cuda:0", seed=0); enh.load_models()
frames = [f.copy() for f in synth.mixed_clip(8, H, W, seed=1)]
dev = torch.from_numpy(np.stack(frames)).cuda()

def t(fn, n=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3

print("device-resident step   %.2f ms" % t(lambda: enh.upscale_device(dev, mode="cnn")))
print("method                 %.2f ms" % t(lambda: enh._process_frame_batch_gpu(frames)))
out = torch.empty((8, 2 * H, 2 * W, 3), dtype=torch.uint8, pin_memory=True)
ptrs = (C.c_void_p * 8)(*[f.ctypes.data for f in frames])
for chunk in (8, 4, 2, 1):
    print("vex_upscale_frames chunk=%d  %.2f ms" % (chunk, t(lambda: enh._lib.vex_upscale_frames(enh._ctx, 1, ptrs, out.data_ptr(), 8, H, W, chunk, 2))))
pin_in = torch.empty((8, H, W, 3), dtype=torch.uint8, pin_memory=True)
pin_in.numpy()[:] = np.stack(frames)
for chunk in (8, 4, 2, 1):
    print("vex_upscale_host   chunk=%d  %.2f ms" % (chunk, t(lambda: enh._lib.vex_upscale_host(enh._ctx, 1, pin_in.data_ptr(), out.data_ptr(), 8, H, W, chunk, 2))))
def alloc():
    x = torch.empty((8, 2 * H, 2 * W, 3), dtype=torch.uint8, pin_memory=True); return x
print("pinned alloc (cached)  %.3f ms" % t(alloc))
st = np.empty((8, H, W, 3), np.uint8)
def cp():
    for i, f in enumerate(frames): st[i] = f
print("host memcpy 8 frames   %.2f ms (%.1f GB/s)" % (t(cp), 8 * H * W * 3 / t(cp) / 1e6))
