#!/usr/bin/env python
"""conv1 (im2col27) at c1 = 256 over a sweep of items per CTA: find the smallest failing size (GPU box)."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "one":
    import numpy as np, torch
    from oracle import upscaler_ref as O
    from video_enhancer_proto_b200 import B200VideoEnhancer, synth
    c1, H, W = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    model = O.make_model(0, widths=(c1, c1, 256, 128))
    enh = B200VideoEnhancer(device="cuda:0", state_dict=O.export_state(model)); enh.load_models()
    f = synth.random_frame(H, W, np.random.default_rng(4))
    x = torch.from_numpy(f).cuda()[None]
    got = enh.cnn_debug_f32(x, tap=1); torch.cuda.synchronize()
    taps = {}
    O.ref_cnn_batch_f32(model, [f], taps) if H * W <= 300000 else None
    err = float((got.cpu() - taps["x1"]).abs().max()) if taps else -1.0
    print("ok", c1, H, W, "items/cta=%.1f" % (H * W / 128 / 148), "err", err, flush=True)
else:
    for c1 in (256, 128):
        for m in (1, 2, 3, 4, 5, 6, 8, 16, 64, 440):
            r = subprocess.run([sys.executable, __file__, "one", str(c1), str(148 * m), "128"], capture_output=True, text=True, timeout=600)
            print((r.stdout.strip().splitlines() or ["FAILED c1=%d m=%d: %s" % (c1, m, r.stderr.strip().splitlines()[-1][:150] if r.stderr.strip() else "?")])[-1], flush=True)
