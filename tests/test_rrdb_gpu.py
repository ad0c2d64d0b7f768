"""GPU parity of the Real-ESRGAN RRDBNet path against the oracle, which is pinned to the reference's
own classes (oracle/make_golden_rrdb.py).

Bars.  The network runs in fp16 with fp32 accumulation; on the CPU, emulating those rounding points gives a maximum error
of 0.5 LSB for the 23-block network (4.8 LSB with bf16), so the north_star CNN bar is reused unchanged for seeded
random-init weights: max-abs <= 2/255 and PSNR >= 50 dB on the fp32 [0,1] output, uint8 within 2 LSB.  Per-layer taps are
compared relative to the layer's range so that a wrong channel slice, a missing residual or a swapped buffer cannot hide
behind the small output range of a random-init network.  The gain-6 models stretch conv_last so that the output leaves
[0,1] on both sides and the clamps and the truncating cast are exercised at full range."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import realesrgan_ref as R
import rrdb_emulation as E
from video_enhancer_proto_b200 import B200RealESRGAN, synth

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAX_ABS, MIN_PSNR = 2.0 / 255.0, 50.0
TAP_REL = 4e-3                       # fp16 activations: 2^-11 per rounding, a few dozen roundings deep


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return float("inf") if mse == 0 else 10.0 * np.log10(1.0 / mse)


def dev(frames):
    return torch.from_numpy(np.stack(frames)).cuda()


def handler(model, **kw):
    h = B200RealESRGAN(device="cuda:0", state_dict=R.export_state(model), **kw)
    h._ensure_model_loaded()
    return h


@pytest.mark.parametrize("scale,nb,seed,gain", [(4, 2, 0, 1.0), (2, 3, 1, 6.0), (4, 1, 3, 6.0)])
@pytest.mark.parametrize("h,w", [(37, 53), (64, 96), (8, 30), (9, 31), (1, 1)])
def test_taps_and_output(scale, nb, seed, gain, h, w):
    model = R.make_network(seed, scale, nb, gain)
    hd = handler(model)
    try:
        rng = np.random.default_rng(100 + h)
        frames = [synth.random_frame(h, w, rng), synth.edges_frame(h, w, rng)]
        d = dev(frames)
        x = torch.from_numpy(np.stack(frames)).permute(0, 3, 1, 2).float() / 255.0
        taps = {}
        with torch.no_grad():
            want = torch.clamp(model(x, taps), 0, 1).numpy()
            # inside the first dense block
            rdb = model.body[0].rdb1
            f0 = taps["first"]
            x1 = rdb.lrelu(rdb.conv1(f0))
            x2 = rdb.lrelu(rdb.conv2(torch.cat((f0, x1), 1)))
            x3 = rdb.lrelu(rdb.conv3(torch.cat((f0, x1, x2), 1)))
            x4 = rdb.lrelu(rdb.conv4(torch.cat((f0, x1, x2, x3), 1)))
            o1 = rdb(f0)
        checks = [(1, taps["first"]), (1001, x1), (1002, x2), (1003, x3), (1004, x4), (1005, o1)]
        checks += [(10 + b, taps[f"body{b}"]) for b in range(nb)]
        checks += [(2, taps["trunk"]), (3, taps["hr"])]
        for tap, ref in checks:
            got = hd.debug_f32(d, tap=tap).cpu().numpy()
            ref = ref.numpy()
            assert got.shape == ref.shape, (tap, got.shape, ref.shape)
            rel = float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-9))
            assert rel <= TAP_REL, (scale, nb, h, w, tap, rel)
        got = hd.debug_f32(d, tap=0).cpu().numpy()
        err = float(np.abs(got - want).max())
        assert err <= MAX_ABS and psnr(got, want) >= MIN_PSNR, (scale, nb, h, w, err, psnr(got, want))
        u8 = hd.upscale_device(d).cpu().numpy()
        for i, f in enumerate(frames):
            ref8 = R.process_single(model, f)
            assert u8[i].shape == ref8.shape
            assert np.abs(u8[i].astype(int) - ref8.astype(int)).max() <= 2, (scale, nb, h, w, i)
    finally:
        hd.close()


def test_golden_fixtures_and_frame_path():
    """The committed vectors minted by the reference's own `_enhance_frame` (BGR in/out, tiled where a side exceeds
    tile_size), through the handler twin's method of the same name."""
    cases = json.load(open(os.path.join(ROOT, "tests", "golden", "rrdb_manifest.json")))["cases"]
    arr = np.load(os.path.join(ROOT, "tests", "golden", "rrdb_cases.npz"))
    for key, c in cases.items():
        model = R.make_network(c["seed"], c["scale"], c["num_block"], c["gain"])
        hd = handler(model, scale=c["scale"], tile_size=c["tile_size"], tile_pad=c["tile_pad"])
        try:
            got = hd._enhance_frame(arr[key + "__in"])
            want = arr[key + "__out_u8"]
            assert got.shape == want.shape and got.dtype == np.uint8
            d = np.abs(got.astype(int) - want.astype(int))
            assert d.max() <= 2, (key, int(d.max()))
            assert (d > 0).mean() < 0.2, (key, float((d > 0).mean()))
            if c["gain"] > 1:
                assert want.max() - want.min() > 100                      # the comparison spans the range
        finally:
            hd.close()


def test_tiled_equals_tiles_run_alone_and_rgb_methods():
    """`_tile_process` (:324-372) on the device: every tile's un-padded core must equal that tile run on its own through
    `_process_single` - bit for bit, the crop and paste are copies - and differ from the un-tiled frame near tile seams
    (10 pixels of context are less than the receptive field), exactly as in the reference."""
    model = R.make_network(4, 2, 2, 6.0)
    hd = handler(model, scale=2, tile_size=64, tile_pad=10)
    try:
        rng = np.random.default_rng(8)
        img = synth.edges_frame(96, 130, rng)
        tiled = hd._tile_process(img)
        assert tiled.shape == (192, 260, 3)
        for (py0, py1, px0, px1, y0, y1, x0, x1) in R.tile_plan(96, 130, 64, 10):
            alone = hd._process_single(np.ascontiguousarray(img[py0:py1, px0:px1]))
            core = alone[(y0 - py0) * 2:(y0 - py0) * 2 + (y1 - y0) * 2, (x0 - px0) * 2:(x0 - px0) * 2 + (x1 - x0) * 2]
            assert np.array_equal(tiled[y0 * 2:y1 * 2, x0 * 2:x1 * 2], core), (y0, x0)
        ref = R.tile_process(model, img, 2, 64, 10)
        assert np.abs(tiled.astype(int) - ref.astype(int)).max() <= 2
        whole = hd._process_single(img)
        assert not np.array_equal(whole, tiled)
        # _enhance_frame == BGR wrapper around the RGB methods (:290, 300)
        bgr = hd._enhance_frame(np.ascontiguousarray(img[:, :, ::-1]))
        assert np.array_equal(bgr[:, :, ::-1], tiled)
    finally:
        hd.close()


@pytest.mark.parametrize("scale,h,w,tile,pad,budget_mb", [(2, 150, 200, 64, 10, 0), (4, 150, 200, 32, 6, 0), (4, 97, 131, 48, 10, 64),
                                                          (2, 128, 192, 64, 0, 0)])
def test_batched_tile_loop_equals_the_sequential_one(scale, h, w, tile, pad, budget_mb):
    """The tiles of a frame run as ONE batch of canvases per layer (csrc/rrdb.cu: rrdb_frame; option rrdb_batch, default
    on), each tile in the corner of a canvas as large as the largest tile with everything outside its own size kept
    zero.  That must be the reference's sequential loop (:342-370) bit for bit: ragged last row / column, more tiles than
    one batch holds (35 > 16), a workspace budget that cuts batches short, and pad 0 (all canvases but the last equal)."""
    model = R.make_network(7, scale, 2, 6.0)
    hd = handler(model, scale=scale, tile_size=tile, tile_pad=pad)
    try:
        img = synth.edges_frame(h, w, np.random.default_rng(h + w))
        if budget_mb:
            hd.set_option("ws_budget_mb", budget_mb)
        n0 = hd.kernel_launches()
        batched = hd._tile_process(img).copy()
        n_batched = hd.kernel_launches() - n0
        prev = hd.set_option("rrdb_batch", 1)                              # one tile after the other
        assert prev > 1
        n0 = hd.kernel_launches()
        sequential = hd._tile_process(img).copy()
        n_seq = hd.kernel_launches() - n0
        assert np.array_equal(batched, sequential)
        assert n_batched < n_seq                                           # it really was batched
        assert int(batched.max()) - int(batched.min()) > 100
        hd.set_option("rrdb_batch", 3)                                     # batches of three, over the stale workspace
        assert np.array_equal(hd._tile_process(img), batched)
        hd.set_option("rrdb_batch", prev)
        assert np.array_equal(hd._tile_process(img), batched)
    finally:
        hd.close()


def test_23_block_network_meets_the_bar():
    """The reference's configuration (num_block 23, scale 4) with seeded random weights: 347 conv layers deep."""
    model = R.make_network(2, 4, 23)
    hd = handler(model)
    try:
        rng = np.random.default_rng(2)
        f = synth.random_frame(40, 56, rng)
        want = R.process_single_f32(model, f)
        got = hd.debug_f32(dev([f]), tap=0).cpu().numpy()[0]
        err = float(np.abs(got - want).max())
        assert err <= MAX_ABS and psnr(got, want) >= MIN_PSNR, (err, psnr(got, want))
        u8 = hd.upscale_device(dev([f])).cpu().numpy()[0]
        assert np.abs(u8.astype(int) - R.process_single(model, f).astype(int)).max() <= 2
        info = hd.get_model_info()
        assert info["parameters"] == sum(p.numel() for p in model.parameters()) == 16697987
        assert info["scale_factor"] == 4 and info["tile_size"] == 512 and "RRDB" in info["architecture"]
    finally:
        hd.close()


def test_batch_and_repeat_are_deterministic():
    model = R.make_network(0, 4, 2, 6.0)
    hd = handler(model)
    try:
        rng = np.random.default_rng(3)
        pool = [synth.random_frame(33, 47, rng), synth.edges_frame(33, 47, rng)]
        a = hd.upscale_device(dev([pool[0], pool[1], pool[0]])).cpu().numpy()
        b = hd.upscale_device(dev([pool[0]])).cpu().numpy()
        c = hd.upscale_device(dev([pool[1]])).cpu().numpy()
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[0]) and np.array_equal(a[1], c[0])
        # a bigger image in between re-lays-out the workspace; the small one must come out the same afterwards
        hd.upscale_device(dev([synth.random_frame(70, 90, rng)]))
        assert np.array_equal(hd.upscale_device(dev([pool[0]])).cpu().numpy()[0], b[0])
    finally:
        hd.close()


def test_restore_video_file_to_file(tmp_path):
    cv2 = pytest.importorskip("cv2")
    src = tmp_path / "in.mp4"
    wr = cv2.VideoWriter(str(src), cv2.VideoWriter_fourcc(*"mp4v"), 24.0, (96, 64))
    if not wr.isOpened():
        pytest.skip("OpenCV cannot write mp4v here")
    rng = np.random.default_rng(0)
    for i in range(7):
        wr.write(synth.demo_frame(i, size=(96, 64), rng=rng, bgr=True))
    wr.release()
    hd = B200RealESRGAN(device="cuda:0", scale=2, seed=0, num_block=1, tile_size=48)
    try:
        s = hd.restore_video(str(src), str(tmp_path / "sub" / "out.mp4"))
        assert s["output_frames"] == s["frames_processed"] == 7 and s["processing_mode"] == "realesrgan"
        assert s["scale_factor"] == 2 and s["input_resolution"] == (64, 96) and s["output_resolution"] == (128, 192)
        cap = cv2.VideoCapture(str(tmp_path / "sub" / "out.mp4"))
        assert int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)) == 192 and int(cap.get(cv2.CAP_PROP_FRAME_COUNT)) == 7
        cap.release()
        with pytest.raises(ValueError, match="Could not open video"):
            hd.restore_video(str(tmp_path / "missing.mp4"), str(tmp_path / "o.mp4"))
    finally:
        hd.close()


def test_full_size_1080p_frame_through_the_tile_loop():
    """The reference configuration at the bench's size: a 1080p frame, 23 blocks, x4, tile 512 / pad 10 (12 tiles -> 4320p).
    The CPU oracle needs minutes for a whole frame, so one INTERIOR tile (532 x 532 with padding) is checked against the
    oracle's `_process_single` on the same padded crop (what `_tile_process` computes for it, :351), and every tile core of
    the frame against the same tile run alone on the GPU (bit for bit)."""
    model = R.make_network(2, 4, 23, 6.0)
    hd = handler(model)
    try:
        rng = np.random.default_rng(5)
        img = synth.edges_frame(1080, 1920, rng)
        out = hd._tile_process(img)
        assert out.shape == (4320, 7680, 3) and out.dtype == np.uint8
        plan = R.tile_plan(1080, 1920, 512, 10)
        assert len(plan) == 12
        for (py0, py1, px0, px1, y0, y1, x0, x1) in (plan[0], plan[5], plan[11]):      # corner, interior, ragged last tile
            alone = hd._process_single(np.ascontiguousarray(img[py0:py1, px0:px1]))
            core = alone[(y0 - py0) * 4:(y0 - py0) * 4 + (y1 - y0) * 4, (x0 - px0) * 4:(x0 - px0) * 4 + (x1 - x0) * 4]
            assert np.array_equal(out[y0 * 4:y1 * 4, x0 * 4:x1 * 4], core), (y0, x0)
        py0, py1, px0, px1, y0, y1, x0, x1 = plan[5]
        ref = R.process_single(model, np.ascontiguousarray(img[py0:py1, px0:px1]))
        want = ref[(y0 - py0) * 4:(y0 - py0) * 4 + (y1 - y0) * 4, (x0 - px0) * 4:(x0 - px0) * 4 + (x1 - x0) * 4]
        got = out[y0 * 4:y1 * 4, x0 * 4:x1 * 4]
        d = np.abs(got.astype(int) - want.astype(int))
        assert want.max() - want.min() > 100                                 # full-range output
        # 23 blocks x gain 6: what the fp16 number format itself costs against fp32 is measured here, on the same tile, by
        # the CPU emulation of the GPU's rounding points (tests/rrdb_emulation.py); the GPU may not be further from the
        # oracle than that, up to the summation-order freedom (1 LSB).  At gain 1 (test_23_block_network_meets_the_bar)
        # the same error is 0.5 LSB and the <= 2 LSB bar holds outright.
        emu = E.process_single(model, np.ascontiguousarray(img[py0:py1, px0:px1]))
        emu = emu[(y0 - py0) * 4:(y0 - py0) * 4 + (y1 - y0) * 4, (x0 - px0) * 4:(x0 - px0) * 4 + (x1 - x0) * 4]
        de = np.abs(emu.astype(int) - want.astype(int))
        dg = np.abs(got.astype(int) - emu.astype(int))

        def psnr(x):
            return 10 * np.log10(255.0 ** 2 / max(float((x.astype(np.float64) ** 2).mean()), 1e-12))

        print(f"[rrdb_1080p] gpu-fp32 max {int(d.max())} mean {d.mean():.4f} psnr {psnr(d):.1f} | emu-fp32 max {int(de.max())} "
              f"mean {de.mean():.4f} psnr {psnr(de):.1f} | gpu-emu max {int(dg.max())} mean {dg.mean():.4f} psnr {psnr(dg):.1f}")
        assert d.max() <= de.max() + 1, (int(d.max()), int(de.max()))
        assert d.mean() <= 1.1 * de.mean() + 0.01, (float(d.mean()), float(de.mean()))
        assert psnr(d) >= psnr(de) - 0.5 and psnr(d) >= 45.0, (psnr(d), psnr(de))
    finally:
        hd.close()
