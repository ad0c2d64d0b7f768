"""Sweep the bicubic strip height (rows per strip) and batch size on the GPU; prints fps and HBM fraction.
Each configuration is a short burst after a 1.5 s pause: back-to-back sustained runs hit the power cap and the
later configurations then read ~12 % slow regardless of their merit."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from video_enhancer_proto_b200 import B200VideoEnhancer

H, W = 1080, 1920
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6483.0
enh = B200VideoEnhancer(device="cuda:0", seed=0)
for n in ((300,) if len(sys.argv) > 1 else (300, 8, 1)):
    x = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device="cuda")
    out = torch.empty((n, 2 * H, 2 * W, 3), dtype=torch.uint8, device="cuda")
    for rows in ([int(a) for a in sys.argv[1:]] or (0, 8, 24, 40, 60, 120, 216, 360)):
        enh.set_option("bicubic_rows", rows)
        torch.cuda.synchronize(); time.sleep(1.5)
        for _ in range(3):
            enh.upscale_device(x, mode="bicubic", out=out)
        reps = 5 if n == 300 else 50
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(2):
            torch.cuda.synchronize(); e0.record()
            for _ in range(reps):
                enh.upscale_device(x, mode="bicubic", out=out)
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / reps)
        gbs = n * H * W * 15 / best / 1e6
        print(f"n={n:4d} rows={rows:4d}  {best:8.4f} ms  {n / best * 1e3:10.0f} fps  {gbs:7.0f} GB/s  frac {gbs / peak:.3f}", flush=True)
