"""CPU: the oracle (oracle/upscaler_ref.py) against the committed golden vectors, which were produced by
running the reference's own code (oracle/make_golden.py AST-extracts GPUVideoEnhancer from app.py)."""
import hashlib

import numpy as np
import pytest

from oracle import upscaler_ref as O
from oracle.make_golden import case_inputs


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_weights_match_reference_init(golden):
    cases, _ = golden
    for seed in (0, 1, 2):
        st = O.export_state(O.make_model(seed))
        assert {k: sha(v) for k, v in st.items()} == cases[f"weights_seed{seed}"]
    st = O.export_state(O.make_stress_model())
    assert {k: sha(v) for k, v in st.items()} == cases["weights_stress"]


@pytest.mark.parametrize("seed,h,w", [(0, 37, 53), (1, 37, 53), (0, 64, 96)])
def test_cnn_oracle_matches_reference_outputs(golden, seed, h, w):
    cases, small = golden
    model = O.make_model(seed)
    for name, frame in case_inputs(h, w, seed * 100 + h).items():
        c = cases[f"cnn_seed{seed}_{h}x{w}_{name}"]
        assert sha(frame) == c["in"]
        assert sha(O.ref_cnn_frame(model, frame)) == c["out_single_rgb"]
        key = f"cnn_seed{seed}_{h}x{w}_{name}__out_u8"
        if key in small.files:
            assert np.array_equal(O.ref_cnn_frame(model, frame), small[key])
    frame = case_inputs(h, w, seed * 100 + h)["random"]
    got = O.ref_cnn_batch(model, [frame, frame[::-1].copy()], out_bgr=True)
    assert [sha(x) for x in got] == cases[f"cnn_seed{seed}_{h}x{w}_random"]["out_batch_bgr"]


def test_cnn_stress_oracle(golden):
    cases, small = golden
    model = O.make_stress_model()
    for name, frame in case_inputs(37, 53, 777 + 37).items():
        c = cases[f"cnn_stress_37x53_{name}"]
        out = O.ref_cnn_frame(model, frame)
        assert sha(out) == c["out_single_rgb"]
    # the stress weights must actually exercise both clamps
    rng_case = cases["cnn_stress_37x53_random"]
    assert rng_case["out_min"] == 0 and rng_case["out_max"] == 255


@pytest.mark.parametrize("h,w", [(1, 1), (2, 3), (5, 7), (37, 53), (64, 96), (360, 480)])
def test_bicubic_oracle_matches_reference_outputs(golden, h, w):
    cases, small = golden
    for name, frame in case_inputs(h, w, 4242 + h).items():
        c = cases[f"bicubic_{h}x{w}_{name}"]
        assert sha(frame) == c["in"]
        out = O.ref_bicubic_frame(frame)
        assert sha(out) == c["out_trunc"]
        key = f"bicubic_{h}x{w}_{name}__out_u8"
        if key in small.files:
            assert np.array_equal(out, small[key])
        # exact-arithmetic restatement (what the CUDA kernel computes) stays within 1 LSB of the reference
        ex = O.np_bicubic2x_u8(frame, 0)
        assert np.abs(ex.astype(int) - out.astype(int)).max() <= 1
        if h >= 2 and w >= 2:
            cv = O.ref_bicubic_cv2(frame)
            assert sha(cv) == c["out_cv2"]
            assert np.abs(O.np_bicubic2x_u8(frame, 1).astype(int) - cv.astype(int)).max() <= 1


def test_bicubic_weights_are_the_keys_kernel():
    # A=-0.75 cubic convolution (ATen upsample_bicubic2d / cv2 INTER_CUBIC) at t=0.25 and 0.75
    A = -0.75
    def k(x):
        x = abs(x)
        if x <= 1:
            return (A + 2) * x**3 - (A + 3) * x**2 + 1
        if x < 2:
            return A * x**3 - 5 * A * x**2 + 8 * A * x - 4 * A
        return 0.0
    for t, wts in ((0.25, O.W_T025), (0.75, O.W_T075)):
        assert np.allclose([k(t + 1), k(t), k(1 - t), k(2 - t)], wts, atol=0, rtol=0)


def test_lanczos4_restatement_matches_opencv():
    """the Q11 tap table and the integer restatement of INTER_LANCZOS4 2x against cv2 itself."""
    cv2 = pytest.importorskip("cv2")
    for t, taps in O.LANCZOS4_TAPS_Q11.items():
        assert O.lanczos4_taps_q11(t) == taps and sum(taps) == 2048
    rng = np.random.default_rng(0)
    for shape in ((1, 1, 3), (5, 7, 3), (20, 33, 3)):
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        want = cv2.resize(img, (shape[1] * 2, shape[0] * 2), interpolation=cv2.INTER_LANCZOS4)
        assert np.array_equal(O.np_lanczos4_2x_u8(img), want), shape


def test_linear_nearest_area_restatements_match_opencv():
    """the remaining BasicUpscalerFallback modes (basic_upscaler_fallback.py:54-56) - the integer
    restatements the CUDA kernel follows against cv2 itself, bit-exact, incl. 1-px-wide / 1-px-high and odd shapes."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    for h, w in ((1, 1), (1, 7), (5, 1), (2, 2), (37, 53), (64, 96), (360, 480)):
        for img in (rng.integers(0, 256, (h, w, 3), dtype=np.uint8), np.full((h, w, 3), 255, np.uint8),
                    (np.indices((h, w)).sum(0) % 2 * 255).astype(np.uint8)[..., None].repeat(3, 2)):
            size = (2 * w, 2 * h)
            assert np.array_equal(O.np_linear2x_u8(img), cv2.resize(img, size, interpolation=cv2.INTER_LINEAR)), (h, w)
            assert np.array_equal(O.np_nearest2x_u8(img), cv2.resize(img, size, interpolation=cv2.INTER_NEAREST)), (h, w)
            assert np.array_equal(O.np_nearest2x_u8(img), cv2.resize(img, size, interpolation=cv2.INTER_AREA)), (h, w)
