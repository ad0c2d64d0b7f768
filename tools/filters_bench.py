#!/usr/bin/env python
"""SURVEY.md §8(f)3 measurement: the BasicUpscalerFallback stages at 1080p -> 2160p on one GPU (frames resident in HBM,
CUDA events, short bursts), algorithmic GB/s against the measured HBM peak, mismatch against OpenCV on one frame, and
OpenCV's own time for the same stage on the host cores."""
import ctypes as C, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cv2, numpy as np, torch
from video_enhancer_proto_b200 import _lib, synth
from video_enhancer_proto_b200.basic_upscaler import B200BasicUpscaler, SHARPEN_KERNEL

H, W, N = 1080, 1920, 64
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6483.0
u = B200BasicUpscaler(device="cuda:0", model_name="LANCZOS")
e = u._enh; e._ensure_ctx()
pool = synth.mixed_clip(4, H, W, seed=3)
x = torch.from_numpy(np.stack([pool[i % 4] for i in range(N)])).cuda()
x2 = torch.empty((N, 2 * H, 2 * W, 3), dtype=torch.uint8, device="cuda")
y1 = torch.empty_like(x); y2 = torch.empty_like(x2)
st = torch.cuda.current_stream().cuda_stream
k9 = (C.c_float * 9)(*[float(v) for v in SHARPEN_KERNEL.ravel()])
stages = {
    "bilateral5 (1080p)": (lambda: e._lib.vex_bilateral5_u8(e._ctx, x.data_ptr(), y1.data_ptr(), N, H, W, 50.0, 50.0, st), 6 * H * W,
                           lambda f: cv2.bilateralFilter(f, 5, 50, 50), lambda: y1[0], pool[0]),
    "lanczos4 2x (1080p->2160p)": (lambda: e._lib.vex_lanczos4_2x_u8(e._ctx, x.data_ptr(), x2.data_ptr(), N, H, W, st), 15 * H * W,
                                   lambda f: cv2.resize(f, (2 * W, 2 * H), interpolation=cv2.INTER_LANCZOS4), lambda: x2[0], pool[0]),
    "linear 2x (1080p->2160p)": (lambda: e._lib.vex_resize2x_u8(e._ctx, x.data_ptr(), x2.data_ptr(), N, H, W, 1, st), 15 * H * W,
                                 lambda f: cv2.resize(f, (2 * W, 2 * H), interpolation=cv2.INTER_LINEAR), lambda: x2[0], pool[0]),
    "nearest/area 2x (1080p->2160p)": (lambda: e._lib.vex_resize2x_u8(e._ctx, x.data_ptr(), x2.data_ptr(), N, H, W, 0, st), 15 * H * W,
                                       lambda f: cv2.resize(f, (2 * W, 2 * H), interpolation=cv2.INTER_NEAREST), lambda: x2[0], pool[0]),
}
out = {}
for name, (fn, bytes_per_frame, cpu, result, src) in stages.items():
    for _ in range(3): _lib.check(e._ctx, fn(), name)
    torch.cuda.synchronize(); time.sleep(1.0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    got = result().cpu().numpy()
    t0 = time.perf_counter(); want = cpu(src); cpu_ms = (time.perf_counter() - t0) * 1e3
    d = np.abs(got.astype(int) - want.astype(int))
    out[name] = {"fps": N / ms * 1e3, "GBps": bytes_per_frame * N / ms / 1e6, "hbm_frac": bytes_per_frame * N / ms / 1e6 / peak,
                 "max_diff": int(d.max()), "mismatch_frac": float((d > 0).mean()), "opencv_cpu_ms_per_frame": cpu_ms,
                 "opencv_threads": cv2.getNumThreads()}
# sharpen runs on the 2160p result of the resize
fn = lambda: e._lib.vex_filter3x3_u8(e._ctx, x2.data_ptr(), y2.data_ptr(), N, 2 * H, 2 * W, k9, st)
for _ in range(3): _lib.check(e._ctx, fn(), "sharpen")
torch.cuda.synchronize(); time.sleep(1.0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): fn()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
src = x2[0].cpu().numpy()
t0 = time.perf_counter(); want = np.clip(cv2.filter2D(src, -1, SHARPEN_KERNEL), 0, 255).astype(np.uint8); cpu_ms = (time.perf_counter() - t0) * 1e3
d = np.abs(y2[0].cpu().numpy().astype(int) - want.astype(int))
out["sharpen 3x3 (2160p)"] = {"fps": N / ms * 1e3, "GBps": 24 * H * W * N / ms / 1e6, "hbm_frac": 24 * H * W * N / ms / 1e6 / peak,
                              "max_diff": int(d.max()), "mismatch_frac": float((d > 0).mean()), "opencv_cpu_ms_per_frame": cpu_ms,
                              "opencv_threads": cv2.getNumThreads()}
# whole LANCZOS pipeline, device-resident
for _ in range(2): u.upscale_device(x)
torch.cuda.synchronize(); time.sleep(1.0)
e0.record()
for _ in range(3): u.upscale_device(x)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
t0 = time.perf_counter()
from oracle import upscaler_ref as O
O.ref_basic_upscaler_frame(pool[0], "LANCZOS"); cpu_ms = (time.perf_counter() - t0) * 1e3
out["LANCZOS pipeline (bilateral + lanczos4 + sharpen)"] = {"fps": N / ms * 1e3, "opencv_cpu_ms_per_frame": cpu_ms, "opencv_threads": cv2.getNumThreads()}
print(json.dumps({"workload": "BasicUpscalerFallback stages, 64 frames 1080p->2160p resident in HBM, one B200", "stages": out}))
