"""GPU: the round-2 host-side boundary - vex_submit / vex_wait tickets, vex_upscale_frames (the plugin-call form),
stream ordering of the shared CNN workspace, and the caller's current device being left alone.
Every result is compared bit for bit with the device-resident entry (vex_cnn2x_u8 / vex_bicubic2x_u8), which the
parity tests pin to the oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

from video_enhancer_proto_b200 import B200VideoEnhancer, _lib, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def enh():
    e = B200VideoEnhancer(device="cuda:0", seed=0)
    assert e.load_models()
    yield e
    e.close()


def dev(frames):
    return torch.from_numpy(np.stack(frames)).cuda()


def pinned(n):
    return torch.empty(n, dtype=torch.uint8).pin_memory()


@pytest.mark.parametrize("mode", ["cnn", "bicubic"])
def test_submit_wait_tickets(enh, mode):
    """Five batches of different sizes submitted back to back, waited for out of order and twice; 10 submits without
    a wait overflow the 8-entry ticket ring (the 9th submit waits for the oldest itself)."""
    rng = np.random.default_rng(1)
    h, w = 40, 56
    mode_id = _lib.VEX_MODE_CNN if mode == "cnn" else _lib.VEX_MODE_BICUBIC
    sizes = [3, 1, 8, 2, 5, 1, 1, 2, 1, 3]
    batches = [[synth.random_frame(h, w, rng) for _ in range(n)] for n in sizes]
    want = [enh.upscale_device(dev(b), mode=mode).cpu().numpy() for b in batches]
    ins = [pinned(n * h * w * 3) for n in sizes]
    outs = [pinned(n * h * w * 12) for n in sizes]
    for b, pi in zip(batches, ins):
        pi.numpy()[:] = np.stack(b).ravel()
    tickets = []
    for n, pi, po in zip(sizes, ins, outs):
        t = C.c_ulonglong()
        rc = enh._lib.vex_submit(enh._ctx, mode_id, pi.data_ptr(), po.data_ptr(), n, h, w, 0, C.byref(t))
        assert rc == 0, enh._lib.vex_last_error(enh._ctx)
        tickets.append(t.value)
    assert tickets == list(range(tickets[0], tickets[0] + len(sizes)))
    for i in (9, 0, 4, 4, 2):                                            # out of order, one twice
        assert enh._lib.vex_wait(enh._ctx, tickets[i]) == 0
    for t in tickets:
        assert enh._lib.vex_wait(enh._ctx, t) == 0
    assert enh._lib.vex_wait(enh._ctx, tickets[-1] + 1) == -1           # never issued
    for n, po, wv in zip(sizes, outs, want):
        assert np.array_equal(po.numpy().reshape(n, 2 * h, 2 * w, 3), wv)


@pytest.mark.parametrize("n", [1, 2, 3, 8, 13])
def test_upscale_frames_equals_device_path(enh, n):
    """`_process_frame_batch_gpu` (app.py:529-553) goes through vex_upscale_frames: separately allocated pageable frames
    in, fresh arrays out, chunks 1, 2, 4, ..., 1.  RGB in -> BGR out.  A non-contiguous frame is accepted too."""
    rng = np.random.default_rng(n)
    h, w = 37, 53
    frames = [synth.random_frame(h, w, rng) for _ in range(n)]
    want = enh.upscale_device(dev(frames), mode="cnn", out_bgr=True).cpu().numpy()
    strided = [f for f in frames]
    strided[0] = np.asfortranarray(frames[0])                            # not C-contiguous
    got = enh._process_frame_batch_gpu(strided)
    assert len(got) == n and all(g.shape == (2 * h, 2 * w, 3) and g.dtype == np.uint8 for g in got)
    for i in range(n):
        assert np.array_equal(got[i], want[i]), i
    again = enh._process_frame_batch_gpu(frames)                         # a second call must not disturb the first result
    for i in range(n):
        assert np.array_equal(got[i], want[i]) and np.array_equal(again[i], want[i])
    with pytest.raises(ValueError):
        enh._process_frame_batch_gpu([])
    with pytest.raises(ValueError):
        enh._process_frame_batch_gpu([frames[0], frames[0][:-1]])


def test_cnn_entries_on_different_streams_are_ordered(enh):
    """vex_cnn2x_u8 runs on the caller's stream, the host-buffer entries on the ctx's private compute
    stream, and both use the ctx's single activation workspace.  No synchronisation between the calls here: the
    workspace event inside cnn_forward must order them."""
    rng = np.random.default_rng(9)
    a = [synth.random_frame(270, 480, rng) for _ in range(8)]           # big enough for the kernels to overlap if unordered
    b = [synth.edges_frame(270, 480, rng) for _ in range(8)]
    da, db = dev(a), dev(b)
    want_a = enh.upscale_device(da, mode="cnn").cpu().numpy()
    want_b = enh.upscale_device(db, mode="cnn").cpu().numpy()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    for _ in range(3):
        out_a = enh.upscale_device(da, mode="cnn")                       # current stream, asynchronous
        got_b = enh._run_host(_lib.VEX_MODE_CNN, b, 0).copy()            # ctx compute stream, no sync in between
        with torch.cuda.stream(side):
            out_a2 = enh.upscale_device(da, mode="cnn")                  # a third stream
        out_b2 = enh.upscale_device(db, mode="cnn")
        torch.cuda.synchronize()
        assert np.array_equal(out_a.cpu().numpy(), want_a)
        assert np.array_equal(got_b, want_b)
        assert np.array_equal(out_a2.cpu().numpy(), want_a)
        assert np.array_equal(out_b2.cpu().numpy(), want_b)


def test_current_device_is_restored():
    """a vex call on ctx->device must not leave that device current in the calling thread."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    e = B200VideoEnhancer(device="cuda:1", seed=0)
    try:
        torch.cuda.set_device(0)
        rng = np.random.default_rng(0)
        f = synth.random_frame(32, 48, rng)
        e.enhance_frame_gpu(f)
        e._enhance_frame_cpu_fallback(f)
        assert torch.cuda.current_device() == 0
        assert torch.empty(1, device="cuda").device.index == 0
    finally:
        e.close()
