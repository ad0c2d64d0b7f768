"""`vex_cnn2x_u8_noalloc` (include/vex.h): the CNN entry point for callers that may not allocate device memory.

It has no counterpart in the reference, so its oracle is the existing entry
point: same kernels in the same order on a different workspace (a static device array inside the library), hence the same
bytes - the bit-identity bar the variant cross-checks of tests/test_cnn_gpu.py use.  Two calls in a row, the second on a
smaller frame and different pixels, so that the first call's left-overs in the shared scratch are there to do harm if they
can.  The file sorts last on purpose: the round's `pytest -x` stops at the first failure."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import upscaler_ref as O
from video_enhancer_proto_b200 import B200VideoEnhancer, _lib, synth

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 176 << 20          # VEX_NOALLOC_SCRATCH_BYTES
VEX_ERR_NOMEM, VEX_ERR_UNSUPPORTED = -5, -4


@pytest.fixture(scope="module")
def enhancer():
    e = B200VideoEnhancer(device="cuda:0", state_dict=O.export_state(O.make_stress_model()))
    assert e.load_models()
    e._ensure_weights()
    yield e
    e.close()


def call(e, name, frames, out, in_bgr=0, out_bgr=0):
    n, h, w, _ = frames.shape
    stream = torch.cuda.current_stream(frames.device).cuda_stream
    return getattr(e._lib, name)(e._ctx, frames.data_ptr(), out.data_ptr(), n, h, w, in_bgr, out_bgr, stream)


def test_two_calls_in_a_row_equal_the_allocating_entry_point(enhancer):
    e = enhancer
    rng = np.random.default_rng(11)
    cases = [(np.stack([synth.edges_frame(120, 168, rng), synth.random_frame(120, 168, rng), synth.checker_frame(120, 168)]), 0, 1),
             (np.stack([synth.random_frame(67, 94, rng)]), 1, 0)]          # smaller, ragged, other channel order
    ins = [torch.from_numpy(c[0]).cuda() for c in cases]
    want = [torch.empty((x.shape[0], 2 * x.shape[1], 2 * x.shape[2], 3), dtype=torch.uint8, device="cuda") for x in ins]
    got = [torch.full_like(t, 0x5A) for t in want]
    for x, o, (_, ib, ob) in zip(ins, want, cases):                     # the existing entry point (also loads every kernel)
        _lib.check(e._ctx, call(e, "vex_cnn2x_u8", x, o, ib, ob), "vex_cnn2x_u8")
    torch.cuda.synchronize()
    launches0 = e.kernel_launches()
    free0 = torch.cuda.mem_get_info()[0]
    for x, o, (_, ib, ob) in zip(ins, got, cases):                      # two calls in a row, nothing in between
        _lib.check(e._ctx, call(e, "vex_cnn2x_u8_noalloc", x, o, ib, ob), "vex_cnn2x_u8_noalloc")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert e.kernel_launches() - launches0 == 2 * 6                      # conv1, conv2-4, conv5, fused tail: the same chain
    for g, wnt in zip(got, want):
        assert wnt.float().std() > 10                                    # a real picture, not a constant
        assert torch.equal(g, wnt)
    # a workspace for these frames would be >= 20 MB; cudaMalloc works in 2 MiB granules
    assert free0 - free1 < (8 << 20), (free0, free1)
    # and a third call after an ordinary one, on the first input again
    again = torch.empty_like(got[0])
    _lib.check(e._ctx, call(e, "vex_cnn2x_u8_noalloc", ins[0], again, cases[0][1], cases[0][2]), "vex_cnn2x_u8_noalloc")
    torch.cuda.synchronize()
    assert torch.equal(again, want[0])


def test_limit_and_refusals(enhancer):
    e = enhancer
    need = C.c_size_t()
    assert e._lib.vex_workspace_bytes(e._ctx, 1, 360, 480, C.byref(need)) == 0
    assert need.value <= SCRATCH_BYTES                                   # the stated limit: one 480 x 360 frame fits
    x = torch.zeros((1, 372, 480, 3), dtype=torch.uint8, device="cuda")
    assert e._lib.vex_workspace_bytes(e._ctx, 1, 372, 480, C.byref(need)) == 0
    assert need.value > SCRATCH_BYTES
    out = torch.full((1, 744, 960, 3), 7, dtype=torch.uint8, device="cuda")
    assert call(e, "vex_cnn2x_u8_noalloc", x, out) == VEX_ERR_NOMEM
    msg = e._lib.vex_last_error(e._ctx).decode()
    assert str(need.value) in msg and str(SCRATCH_BYTES) in msg, msg
    torch.cuda.synchronize()
    assert int(out.min()) == 7 and int(out.max()) == 7                   # nothing ran
    small = torch.zeros((1, 32, 32, 3), dtype=torch.uint8, device="cuda")
    sout = torch.empty((1, 64, 64, 3), dtype=torch.uint8, device="cuda")
    assert e.set_option("profile", 1) == 0
    try:
        assert call(e, "vex_cnn2x_u8_noalloc", small, sout) == VEX_ERR_UNSUPPORTED
    finally:
        e.set_option("profile", 0)
    assert call(e, "vex_cnn2x_u8_noalloc", small, sout) == 0
    torch.cuda.synchronize()


def test_one_whole_480x360_frame_at_the_limit(enhancer):
    e = enhancer
    x = torch.from_numpy(synth.demo_frame(3)[None]).cuda()
    assert tuple(x.shape) == (1, 360, 480, 3)
    want = e.upscale_device(x)
    got = torch.empty_like(want)
    _lib.check(e._ctx, call(e, "vex_cnn2x_u8_noalloc", x, got), "vex_cnn2x_u8_noalloc")
    torch.cuda.synchronize()
    assert torch.equal(got, want)
