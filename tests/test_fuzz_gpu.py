"""Seeded random-shape sweeps: every kernel family against its checker on shapes nobody picked by hand (odd sizes, widths
around the 16 / 30 / 32 / 128-pixel tile edges, single rows and columns, ragged batches).  Small shapes, so the CPU side
stays in seconds."""
import numpy as np
import pytest
import torch

cv2 = pytest.importorskip("cv2")

from oracle import upscaler_ref as O
from video_enhancer_proto_b200 import B200VideoEnhancer, synth
from video_enhancer_proto_b200.basic_upscaler import B200BasicUpscaler

pytestmark = pytest.mark.gpu


def shapes(seed, count, hmax, wmax, specials=()):
    rng = np.random.default_rng(seed)
    out = [(int(rng.integers(1, hmax + 1)), int(rng.integers(1, wmax + 1))) for _ in range(count)]
    return list(specials) + out


def test_bicubic_random_shapes_bit_exact(enhancer):
    """Both rounding modes, strip kernel (W % 16 == 0) and generic kernel, against the exact-arithmetic restatement."""
    rng = np.random.default_rng(100)
    for h, w in shapes(1, 14, 70, 150, specials=[(1, 16), (3, 128), (9, 144), (2, 272), (121, 16), (7, 1)]):
        n = int(rng.integers(1, 4))
        frames = [synth.random_frame(h, w, rng) if i % 2 == 0 else synth.edges_frame(h, w, rng) for i in range(n)]
        t = torch.from_numpy(np.stack(frames)).cuda()
        for mode in (0, 1):
            got = enhancer.upscale_device(t, mode="bicubic", round_mode=mode).cpu().numpy()
            for i, f in enumerate(frames):
                assert np.array_equal(got[i], O.np_bicubic2x_u8(f, mode)), (h, w, mode, i)


def test_cnn_random_shapes_meet_the_bar():
    """Fused-tail and conv tiles with ragged edges in both directions: fp32 output against the fp32 reference."""
    model = O.make_model(5)
    enh = B200VideoEnhancer(device="cuda:0", state_dict=O.export_state(model))
    try:
        rng = np.random.default_rng(200)
        for h, w in shapes(2, 7, 50, 75, specials=[(1, 31), (31, 1), (6, 30), (7, 31), (3, 33), (17, 15)]):
            n = int(rng.integers(1, 4))
            frames = [synth.random_frame(h, w, rng) for _ in range(n)]
            want = O.ref_cnn_batch_f32(model, frames).numpy()
            got = enh.cnn_debug_f32(torch.from_numpy(np.stack(frames)).cuda(), tap=0).cpu().numpy()
            err = float(np.abs(got - want).max())
            rel = err / max(float(np.abs(want).max()), 1e-12)
            assert err <= 2.0 / 255.0 and rel <= 5e-2 and O.psnr(got, want) >= 50.0, (h, w, n, err, rel)
            u8 = enh.upscale_device(torch.from_numpy(np.stack(frames)).cuda(), mode="cnn").cpu().numpy()
            ref8 = np.stack([O.quantize_u8(torch.from_numpy(want[i])) for i in range(n)])
            assert np.abs(u8.astype(int) - ref8.astype(int)).max() <= 2, (h, w)
    finally:
        enh.close()


def test_filter_stages_random_shapes():
    """Lanczos4 bit-exact, bilateral and filter2D within 1 LSB of OpenCV, including widths that take the 16-byte paths."""
    up = B200BasicUpscaler(device="cuda:0", model_name="LANCZOS")
    try:
        rng = np.random.default_rng(300)
        for h, w in shapes(3, 10, 60, 100, specials=[(2, 16), (5, 32), (33, 48), (1, 7), (40, 8)]):
            f = synth.random_frame(h, w, rng) if (h + w) % 2 else synth.edges_frame(h, w, rng)
            den_ref = cv2.bilateralFilter(f, 5, 50, 50)
            got = up.enhance_frames([f])[0]
            want = O.ref_basic_upscaler_frame(f, "LANCZOS")
            d = np.abs(got.astype(int) - want.astype(int))
            assert got.shape == (2 * h, 2 * w, 3) and d.max() <= 2, (h, w, int(d.max()))
            x = torch.from_numpy(f[None]).cuda()
            lz = up.upscale_device(x, sharpen=False).cpu().numpy()[0]            # bilateral + Lanczos4, no sharpen
            ref = cv2.resize(den_ref, (2 * w, 2 * h), interpolation=cv2.INTER_LANCZOS4)
            assert np.abs(lz.astype(int) - ref.astype(int)).max() <= 2, (h, w)
    finally:
        up.close()


def test_resize2x_random_shapes_bit_exact():
    """LINEAR / NEAREST / AREA (basic_upscaler_fallback.py:54-56) on random shapes and batch sizes against the integer
    restatements and cv2: widths with w % 4 == 0 take the packed word kernel, the rest the byte kernel (even / odd widths)."""
    up = B200BasicUpscaler(device="cuda:0", model_name="LINEAR")
    try:
        rng = np.random.default_rng(400)
        for h, w in shapes(4, 16, 70, 160, specials=[(1, 4), (1, 1), (3, 8), (70, 4), (11, 132), (2, 1024), (9, 6), (9, 7)]):
            n = int(rng.integers(1, 4))
            frames = [synth.random_frame(h, w, rng) if i % 2 == 0 else synth.edges_frame(h, w, rng) for i in range(n)]
            x = torch.from_numpy(np.stack(frames)).cuda()
            for name, want_fn, flag in (("LINEAR", O.np_linear2x_u8, cv2.INTER_LINEAR), ("NEAREST", O.np_nearest2x_u8, cv2.INTER_NEAREST),
                                        ("AREA", O.np_nearest2x_u8, cv2.INTER_AREA)):
                up.method = name
                got = up.upscale_device(x, sharpen=False).cpu().numpy()
                for i, f in enumerate(frames):
                    assert np.array_equal(got[i], want_fn(f)), (name, h, w, i)
                    assert np.array_equal(got[i], cv2.resize(f, (2 * w, 2 * h), interpolation=flag)), (name, h, w, i)
    finally:
        up.close()
