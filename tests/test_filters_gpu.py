"""the BasicUpscalerFallback stages (Lanczos4 2x, 5x5 bilateral, 3x3 sharpen) on the GPU against
OpenCV itself, the library the reference calls (basic_upscaler_fallback.py:113-128).

Bars: Lanczos4 is integer fixed-point -> bit-exact.  filter2D and bilateralFilter accumulate in fp32 (OpenCV's SIMD path
uses fused multiply-adds whose grouping is not part of its contract) -> |diff| <= 1 LSB with a small mismatch fraction."""
import numpy as np
import pytest
import torch

cv2 = pytest.importorskip("cv2")

from oracle import upscaler_ref as O
from video_enhancer_proto_b200 import _lib, synth
from video_enhancer_proto_b200.basic_upscaler import B200BasicUpscaler, SHARPEN_KERNEL

pytestmark = pytest.mark.gpu
SHAPES = [(1, 1), (2, 3), (9, 11), (37, 53), (64, 96), (130, 70), (360, 480)]


@pytest.fixture(scope="module")
def lanczos():
    u = B200BasicUpscaler(device="cuda:0", model_name="LANCZOS")
    yield u
    u.close()


@pytest.fixture(scope="module")
def cubic():
    u = B200BasicUpscaler(device="cuda:0", model_name="CUBIC")
    yield u
    u.close()


def frames_for(h, w, seed=0):
    rng = np.random.default_rng(seed)
    return [synth.random_frame(h, w, rng), synth.edges_frame(h, w, rng), np.zeros((h, w, 3), np.uint8),
            np.full((h, w, 3), 255, np.uint8)]


def call(u, name, x, *args, out_scale=1):
    e = u._enh
    e._ensure_ctx()
    n, h, w, _ = x.shape
    out = torch.empty((n, out_scale * h, out_scale * w, 3), dtype=torch.uint8, device="cuda")
    rc = getattr(e._lib, name)(e._ctx, x.data_ptr(), out.data_ptr(), n, h, w, *args, torch.cuda.current_stream().cuda_stream)
    _lib.check(e._ctx, rc, name)
    return out.cpu().numpy()


@pytest.mark.parametrize("h,w", SHAPES)
def test_lanczos4_bit_exact(lanczos, h, w):
    fr = frames_for(h, w)
    got = call(lanczos, "vex_lanczos4_2x_u8", torch.from_numpy(np.stack(fr)).cuda(), out_scale=2)
    for i, f in enumerate(fr):
        want = cv2.resize(f, (2 * w, 2 * h), interpolation=cv2.INTER_LANCZOS4)
        assert np.array_equal(got[i], want), (h, w, i)


def test_lanczos4_extreme_overshoot(lanczos):
    """1-px checkerboards and stripes drive the fixed-point sums to their extremes (saturation on both sides)."""
    h, w = 40, 56
    yy, xx = np.mgrid[0:h, 0:w]
    pats = [((yy + xx) & 1) * 255, (xx & 1) * 255, (yy & 1) * 255, ((xx // 2 + yy // 3) & 1) * 255]
    fr = [np.repeat(p[:, :, None], 3, axis=2).astype(np.uint8) for p in pats]
    got = call(lanczos, "vex_lanczos4_2x_u8", torch.from_numpy(np.stack(fr)).cuda(), out_scale=2)
    for i, f in enumerate(fr):
        assert np.array_equal(got[i], cv2.resize(f, (2 * w, 2 * h), interpolation=cv2.INTER_LANCZOS4)), i


@pytest.mark.parametrize("h,w", SHAPES)
def test_sharpen_filter2d(cubic, h, w):
    import ctypes as C
    fr = frames_for(h, w, seed=1)
    k9 = (C.c_float * 9)(*[float(v) for v in SHARPEN_KERNEL.ravel()])
    got = call(cubic, "vex_filter3x3_u8", torch.from_numpy(np.stack(fr)).cuda(), k9)
    ident = (C.c_float * 9)(0, 0, 0, 0, 1, 0, 0, 0, 0)
    same = call(cubic, "vex_filter3x3_u8", torch.from_numpy(np.stack(fr)).cuda(), ident)
    for i, f in enumerate(fr):
        want = np.clip(cv2.filter2D(f, -1, SHARPEN_KERNEL), 0, 255).astype(np.uint8)
        d = np.abs(got[i].astype(int) - want.astype(int))
        assert d.max() <= 1 and (d > 0).mean() <= 0.02, (h, w, i, int(d.max()), float((d > 0).mean()))
        assert np.array_equal(same[i], f)


@pytest.mark.parametrize("h,w", SHAPES)
def test_bilateral(lanczos, h, w):
    fr = frames_for(h, w, seed=2)
    got = call(lanczos, "vex_bilateral5_u8", torch.from_numpy(np.stack(fr)).cuda(), 50.0, 50.0)
    for i, f in enumerate(fr):
        want = cv2.bilateralFilter(f, 5, 50, 50)
        d = np.abs(got[i].astype(int) - want.astype(int))
        assert d.max() <= 1 and (d > 0).mean() <= 0.02, (h, w, i, int(d.max()), float((d > 0).mean()))


@pytest.mark.parametrize("name", ["LANCZOS", "CUBIC"])
def test_whole_frame_pipeline_matches_reference_body(lanczos, cubic, name):
    """The three stages chained exactly as basic_upscaler_fallback.py:113-128 does, against the oracle's verbatim body."""
    u = lanczos if name == "LANCZOS" else cubic
    rng = np.random.default_rng(5)
    fr = [synth.demo_frame(3, size=(96, 64), rng=rng, bgr=True), synth.random_frame(64, 96, rng), synth.edges_frame(64, 96, rng)]
    got = u.enhance_frames(fr)
    for g, f in zip(got, fr):
        want = O.ref_basic_upscaler_frame(f, name)
        d = np.abs(g.astype(int) - want.astype(int))
        # a 1-LSB difference before the sharpening filter moves its output by at most 0.9 + 8*0.1 = 1.7 -> 2 LSB
        assert d.max() <= 2 and (d > 0).mean() <= 0.05, (name, int(d.max()), float((d > 0).mean()))
        assert g.shape == (128, 192, 3)


CV_FLAGS = {"NEAREST": cv2.INTER_NEAREST, "LINEAR": cv2.INTER_LINEAR, "CUBIC": cv2.INTER_CUBIC, "AREA": cv2.INTER_AREA,
            "LANCZOS": cv2.INTER_LANCZOS4}


@pytest.mark.parametrize("h,w", SHAPES + [(5, 1), (1, 7), (33, 64)])
def test_resize2x_linear_nearest_area_bit_exact(cubic, h, w):
    """vex_resize2x_u8 with OpenCV's own flag values against cv2.resize (basic_upscaler_fallback.py:51-57,120) and the
    integer restatements: NEAREST / LINEAR / AREA / LANCZOS4 bit-exact, CUBIC within 1 LSB; odd widths take the
    byte-store kernel, even widths the word-store one."""
    fr = frames_for(h, w, seed=h * 131 + w) + [(np.indices((h, w)).sum(0) % 2 * 255).astype(np.uint8)[..., None].repeat(3, 2)]
    x = torch.from_numpy(np.stack(fr)).cuda()
    for name, flag in CV_FLAGS.items():
        assert int(flag) == B200BasicUpscaler.CV_INTERPOLATION[name]
        got = call(cubic, "vex_resize2x_u8", x, int(flag), out_scale=2)
        for g, f in zip(got, fr):
            want = cv2.resize(f, (2 * w, 2 * h), interpolation=flag)
            if name == "CUBIC":
                assert np.abs(g.astype(int) - want.astype(int)).max() <= 1, (name, h, w)
            else:
                assert np.array_equal(g, want), (name, h, w)
            if name == "LINEAR":
                assert np.array_equal(g, O.np_linear2x_u8(f))
            if name in ("NEAREST", "AREA"):
                assert np.array_equal(g, O.np_nearest2x_u8(f))


def test_resize2x_rejects_unknown_flag_and_unaligned_output_is_correct(cubic):
    e = cubic._enh
    e._ensure_ctx()
    x = torch.zeros((1, 4, 4, 3), dtype=torch.uint8, device="cuda")
    y = torch.empty((1, 8, 8, 3), dtype=torch.uint8, device="cuda")
    rc = e._lib.vex_resize2x_u8(e._ctx, x.data_ptr(), y.data_ptr(), 1, 4, 4, 7, torch.cuda.current_stream().cuda_stream)
    assert rc < 0 and "interpolation" in e._lib.vex_last_error(e._ctx).decode()
    # an output buffer that is not word-aligned (a view one byte into an allocation) must take the byte-store path
    rng = np.random.default_rng(3)
    f = synth.random_frame(6, 8, rng)
    xin = torch.from_numpy(f[None]).cuda()
    buf = torch.zeros(12 * 16 * 3 + 8, dtype=torch.uint8, device="cuda")
    rc = e._lib.vex_resize2x_u8(e._ctx, xin.data_ptr(), buf.data_ptr() + 1, 1, 6, 8, 1, torch.cuda.current_stream().cuda_stream)
    _lib.check(e._ctx, rc, "vex_resize2x_u8")
    got = buf[1:1 + 12 * 16 * 3].cpu().numpy().reshape(12, 16, 3)
    assert np.array_equal(got, cv2.resize(f, (16, 12), interpolation=cv2.INTER_LINEAR))
    assert int(buf[0]) == 0 and int(buf[1 + 12 * 16 * 3:].max()) == 0


@pytest.mark.parametrize("name", ["LINEAR", "AREA", "NEAREST"])
def test_remaining_modes_pipeline_matches_reference_body(name):
    """The whole per-frame body of the reference for the three modes added last (resize -> 0.1-scaled sharpen -> clip)."""
    u = B200BasicUpscaler(device="cuda:0", model_name=name)
    try:
        rng = np.random.default_rng(11)
        fr = [synth.demo_frame(3, size=(96, 64), rng=rng, bgr=True), synth.random_frame(64, 96, rng), synth.edges_frame(64, 96, rng)]
        for g, f in zip(u.enhance_frames(fr), fr):
            want = O.ref_basic_upscaler_frame(f, name)
            d = np.abs(g.astype(int) - want.astype(int))
            assert g.shape == (128, 192, 3) and d.max() <= 1 and (d > 0).mean() <= 0.01, (name, int(d.max()))
    finally:
        u.close()


def test_unknown_model_name_is_lanczos_without_the_prefilter():
    """an unknown name falls back to INTER_LANCZOS4 for the resize only (dict.get default,
    basic_upscaler_fallback.py:59-62); the bilateral pre-filter is gated on the NAME being 'LANCZOS' (:115-117), so it
    must not run.  Reference body for such a name = resize LANCZOS4 -> 0.1-scaled sharpen -> clip."""
    u = B200BasicUpscaler(device="cuda:0", model_name="something-else")
    try:
        assert u.method == "LANCZOS" and not u._denoise
        rng = np.random.default_rng(12)
        fr = [synth.random_frame(64, 96, rng), synth.edges_frame(64, 96, rng)]
        kernel = np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]], dtype=np.float32) * 0.1
        for g, f in zip(u.enhance_frames(fr), fr):
            up = cv2.resize(f, (192, 128), interpolation=cv2.INTER_LANCZOS4)
            want = np.clip(cv2.filter2D(up, -1, kernel), 0, 255).astype(np.uint8)
            with_prefilter = O.ref_basic_upscaler_frame(f, "LANCZOS")
            d = np.abs(g.astype(int) - want.astype(int))
            assert d.max() <= 1 and (d > 0).mean() <= 0.01, int(d.max())
            assert np.abs(g.astype(int) - with_prefilter.astype(int)).mean() > np.abs(d).mean()
    finally:
        u.close()


@pytest.mark.parametrize("h,w", [(1, 4), (2, 8), (17, 12), (9, 516), (64, 96)])
def test_resize2x_word_kernel_matches_byte_kernel(cubic, h, w):
    """w % 4 == 0 takes the packed 4-pixel kernel (one unit = first and last at w = 4); option force_generic_resize runs
    the byte-indexed kernel on the same input: identical, and both equal cv2."""
    e = cubic._enh
    e._ensure_ctx()
    rng = np.random.default_rng(h * 1000 + w)
    fr = [synth.random_frame(h, w, rng), synth.edges_frame(h, w, rng), np.full((h, w, 3), 255, np.uint8)]
    x = torch.from_numpy(np.stack(fr)).cuda()
    for flag in (cv2.INTER_LINEAR, cv2.INTER_NEAREST, cv2.INTER_AREA):
        fast = call(cubic, "vex_resize2x_u8", x, int(flag), out_scale=2)
        prev = e._lib.vex_set_option(e._ctx, b"force_generic_resize", 1)
        try:
            slow = call(cubic, "vex_resize2x_u8", x, int(flag), out_scale=2)
        finally:
            e._lib.vex_set_option(e._ctx, b"force_generic_resize", max(prev, 0))
        assert np.array_equal(fast, slow), (h, w, flag)
        for g, f in zip(fast, fr):
            assert np.array_equal(g, cv2.resize(f, (2 * w, 2 * h), interpolation=flag)), (h, w, flag)


def test_resize2x_1080p_full_size(cubic):
    rng = np.random.default_rng(9)
    f = synth.edges_frame(1080, 1920, rng)
    # 5 frames: enough threads for the 16-row strips of the word kernel (smaller batches take 8-row strips)
    x = torch.from_numpy(np.stack([f, synth.random_frame(1080, 1920, rng), f[::-1].copy(), f[:, ::-1].copy(), 255 - f])).cuda()
    for flag in (cv2.INTER_LINEAR, cv2.INTER_NEAREST, cv2.INTER_AREA):
        got = call(cubic, "vex_resize2x_u8", x, int(flag), out_scale=2)
        for i in range(5):
            assert np.array_equal(got[i], cv2.resize(x[i].cpu().numpy(), (3840, 2160), interpolation=flag))


def test_1080p_full_size(lanczos):
    rng = np.random.default_rng(7)
    f = synth.edges_frame(1080, 1920, rng)
    x = torch.from_numpy(f[None]).cuda()
    got = call(lanczos, "vex_lanczos4_2x_u8", x, out_scale=2)[0]
    assert np.array_equal(got, cv2.resize(f, (3840, 2160), interpolation=cv2.INTER_LANCZOS4))
    den = call(lanczos, "vex_bilateral5_u8", x, 50.0, 50.0)[0]
    d = np.abs(den.astype(int) - cv2.bilateralFilter(f, 5, 50, 50).astype(int))
    assert d.max() <= 1 and (d > 0).mean() <= 0.02


def test_video_and_info(tmp_path, cubic):
    src = tmp_path / "in.mp4"
    wr = cv2.VideoWriter(str(src), cv2.VideoWriter_fourcc(*"mp4v"), 25.0, (64, 48))
    if not wr.isOpened():
        pytest.skip("OpenCV cannot write mp4v here")
    rng = np.random.default_rng(0)
    for i in range(11):
        wr.write(synth.demo_frame(i, size=(64, 48), rng=rng, bgr=True))
    wr.release()
    st = cubic.enhance_video(str(src), str(tmp_path / "o" / "out.mp4"))
    assert st["frames_processed"] == 11 and st["frames_failed"] == 0 and st["model"] == "BasicUpscaler-CUBIC"
    assert st["input_resolution"] == "64x48" and st["output_resolution"] == "128x96" and st["interpolation_method"] == "CUBIC"
    cap = cv2.VideoCapture(str(tmp_path / "o" / "out.mp4"))
    assert int(cap.get(cv2.CAP_PROP_FRAME_COUNT)) == 11 and int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)) == 128
    cap.release()
    info = cubic.get_model_info()
    assert info["name"] == "BasicUpscaler" and info["interpolation_method"] == "CUBIC" and info["scale"] == 2
    with pytest.raises(ValueError, match="Could not open video"):
        cubic.enhance_video(str(tmp_path / "missing.mp4"), str(tmp_path / "x.mp4"))
    with pytest.raises(ValueError):
        B200BasicUpscaler(model_name="NEAREST", scale=4)
