"""GPU parity: the bicubic CUDA path (through the C ABI) against the oracle.

Bars (BASELINE.json north_star): |out - reference torch path| <= 1 LSB everywhere.  Stronger, checked too:
the kernel equals the EXACT-arithmetic restatement (oracle.np_bicubic2x_u8) bit for bit, in both
rounding modes, which pins every tap index, weight, border clamp and the truncation rule.
"""
import numpy as np
import pytest
import torch

from oracle import upscaler_ref as O
from oracle.make_golden import case_inputs
from video_enhancer_proto_b200 import synth

pytestmark = pytest.mark.gpu

SHAPES = [(1, 1), (2, 3), (5, 7), (37, 53), (64, 96), (33, 128), (40, 144), (96, 256), (360, 480), (480, 640)]


def run(enh, frames, round_mode=0):
    t = torch.from_numpy(np.stack(frames)).cuda()
    out = enh.upscale_device(t, mode="bicubic", round_mode=round_mode)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("h,w", SHAPES)
def test_bicubic_matches_oracle(enhancer, h, w):
    inputs = case_inputs(h, w, 4242 + h)
    names = list(inputs)
    got = run(enhancer, [inputs[k] for k in names], 0)
    got_r = run(enhancer, [inputs[k] for k in names], 1)
    for i, k in enumerate(names):
        ref = O.ref_bicubic_frame(inputs[k])                     # reference torch path (app.py:364-376)
        d = np.abs(got[i].astype(int) - ref.astype(int))
        assert d.max() <= 1, (k, int(d.max()))
        assert np.array_equal(got[i], O.np_bicubic2x_u8(inputs[k], 0)), k     # exact arithmetic, truncation
        assert np.array_equal(got_r[i], O.np_bicubic2x_u8(inputs[k], 1)), k   # exact arithmetic, rounding
        if h >= 2 and w >= 2:
            cv = O.ref_bicubic_cv2(inputs[k])                    # OpenCV twin (basic_upscaler_fallback.py:142)
            assert np.abs(got_r[i].astype(int) - cv.astype(int)).max() <= 1, k


def test_bicubic_golden_fixtures(enhancer, golden):
    cases, small = golden
    for key in small.files:
        if key.startswith("bicubic_") and key.endswith("__in"):
            frame = small[key]
            want = small[key.replace("__in", "__out_u8")]        # produced by the reference's own method
            got = run(enhancer, [frame], 0)[0]
            assert np.abs(got.astype(int) - want.astype(int)).max() <= 1, key


def test_generic_and_tile_kernels_agree(enhancer):
    rng = np.random.default_rng(3)
    frames = [synth.random_frame(70, 272, rng), synth.edges_frame(70, 272, rng)]
    a = run(enhancer, frames, 0)
    prev = enhancer.set_option("force_generic_bicubic", 1)
    try:
        b = run(enhancer, frames, 0)
    finally:
        enhancer.set_option("force_generic_bicubic", prev)
    assert np.array_equal(a, b)


def test_bicubic_1080p_full_size(enhancer):
    """BASELINE.json configs[1] shape.  Full-size checks: exact restatement on whole frames, +-1 LSB vs the
    reference path, and size-independent properties (constant frames are fixed points; flips commute)."""
    frames = synth.mixed_clip(3, 1080, 1920, seed=1234)
    got = run(enhancer, frames, 0)
    assert got.shape == (3, 2160, 3840, 3)
    for i, f in enumerate(frames):
        assert np.array_equal(got[i], O.np_bicubic2x_u8(f, 0)), i
    ref = O.ref_bicubic_frame(frames[0])
    assert np.abs(got[0].astype(int) - ref.astype(int)).max() <= 1
    for v in (0, 1, 77, 254, 255):
        c = run(enhancer, [synth.constant_frame(1080, 1920, v)], 0)[0]
        assert c.min() == v and c.max() == v                      # weights sum to 1 exactly
    flip = run(enhancer, [np.ascontiguousarray(frames[0][::-1, ::-1])], 0)[0]
    assert np.array_equal(flip[::-1, ::-1], got[0])               # the kernel is symmetric under 180-degree rotation


def test_bicubic_host_entry_and_frame_methods(enhancer):
    rng = np.random.default_rng(11)
    frames = [synth.random_frame(64, 96, rng) for _ in range(5)]
    enhancer.batch_size = 2                                       # forces 3 chunks over the 2 streaming slots
    try:
        from video_enhancer_proto_b200 import _lib
        out = enhancer._run_host(_lib.VEX_MODE_BICUBIC, frames, 0).copy()
    finally:
        enhancer.batch_size = 8
    for i, f in enumerate(frames):
        assert np.array_equal(out[i], O.np_bicubic2x_u8(f, 0))
    one = enhancer._enhance_frame_cpu_fallback(frames[0])         # reference method name, GPU arithmetic
    assert one.shape == (128, 192, 3) and np.array_equal(one, out[0])


def test_bad_arguments_are_errors_not_crashes(enhancer):
    from video_enhancer_proto_b200 import _lib
    lib = _lib.load()
    assert lib.vex_bicubic2x_u8(enhancer._ctx, None, None, 1, 4, 4, 0, None) == -1
    t = torch.zeros((1, 4, 4, 3), dtype=torch.uint8, device="cuda")
    o = torch.zeros((1, 8, 8, 3), dtype=torch.uint8, device="cuda")
    assert lib.vex_bicubic2x_u8(enhancer._ctx, t.data_ptr(), o.data_ptr(), 1, 4, 4, 7, None) == -1
    assert b"bad argument" in lib.vex_last_error(enhancer._ctx)
    assert lib.vex_bicubic2x_u8(enhancer._ctx, t.data_ptr(), o.data_ptr(), 0, 4, 4, 0, None) == -1


def test_configs1_all_300_frames_within_one_lsb_of_the_torch_path(enhancer):
    """BASELINE.md section 3 / VERDICT r1 weak #1c: bicubic within +-1 LSB of the reference torch-CPU path
    (`_enhance_frame_cpu_fallback`, app.py:364-376) on ALL 300 frames of configs[1] (100 random + 100 demo pattern +
    100 edges / gradients, seed 1234, 1080p), and bit-equal to the exact restatement on every 25th frame."""
    import torch
    from oracle import upscaler_ref as O
    from video_enhancer_proto_b200 import synth
    torch.set_num_threads(max(1, (__import__("os").cpu_count() or 2)))
    frames = synth.mixed_clip(300, 1080, 1920, seed=1234)
    worst, differing = 0, 0
    for f0 in range(0, 300, 20):
        chunk = frames[f0:f0 + 20]
        got = enhancer.upscale_device(torch.from_numpy(np.stack(chunk)).cuda(), mode="bicubic").cpu().numpy()
        for i, f in enumerate(chunk):
            ref = O.ref_bicubic_frame(f)
            d = np.abs(got[i].astype(np.int16) - ref.astype(np.int16))
            worst = max(worst, int(d.max()))
            differing += int((d > 0).sum())
            if (f0 + i) % 25 == 0:
                assert np.array_equal(got[i], O.np_bicubic2x_u8(f, 0)), f0 + i
    assert worst <= 1, worst                                      # the north_star bar
    # where they differ: the reference computes (v / 255) interpolated x 255 in fp32, which lands a hair below an integer
    # that exact arithmetic reaches (flat and gradient regions of the pattern frames: 1.7 % of all bytes), then truncates
    assert differing / (300 * 2160 * 3840 * 3) < 0.05
