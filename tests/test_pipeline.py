"""Host-side video pipeline: ordering, slot rotation, back-pressure and error propagation of
the reader -> GPU -> writer threads.  No CUDA here; the GPU stage is a stand-in callback."""
import threading
import time

import numpy as np
import pytest

from video_enhancer_proto_b200 import pipeline


def make_reader(n_batches, delay=0.0):
    it = iter(range(n_batches))

    def read():
        if delay:
            time.sleep(delay)
        return next(it, None)
    return read


def test_order_count_and_threads():
    seen, threads = [], set()

    def process(b, slot):
        threads.add(("p", threading.current_thread().name))
        return b * 10

    def write(r):
        threads.add(("w", threading.current_thread().name))
        seen.append(r)

    n = pipeline.run_ordered_pipeline(make_reader(37), process, write, depth=2)
    assert n == 37 and seen == [10 * i for i in range(37)]
    assert ("w", "vex-writer") in threads and ("p", threading.current_thread().name) in threads


def test_empty_stream():
    assert pipeline.run_ordered_pipeline(lambda: None, lambda b, s: b, lambda r: None) == 0


@pytest.mark.parametrize("depth", [1, 2, 4])
def test_slot_not_reused_while_writer_holds_it(depth):
    """Results alias per-slot buffers (the pinned output buffers of the real path): with a slow writer the producer
    must block instead of overwriting a slot that is still queued or being written."""
    bufs = [np.zeros(4, np.int64) for _ in range(pipeline.n_slots(depth))]
    got = []

    def process(b, slot):
        bufs[slot][:] = b
        return bufs[slot]

    def write(r):
        time.sleep(0.003)
        got.append(int(r[0]))
        assert (r == r[0]).all()

    pipeline.run_ordered_pipeline(make_reader(40), process, write, depth=depth)
    assert got == list(range(40))


def test_slow_reader_and_fast_writer():
    got = []
    pipeline.run_ordered_pipeline(make_reader(12, delay=0.002), lambda b, s: b, got.append, depth=2)
    assert got == list(range(12))


@pytest.mark.parametrize("stage", ["read", "process", "write"])
def test_errors_propagate_and_nothing_hangs(stage):
    it = iter(range(1000))

    def read():
        v = next(it)
        if stage == "read" and v == 5:
            raise OSError("decode failed")
        return v

    def process(b, slot):
        if stage == "process" and b == 5:
            raise RuntimeError("vex_upscale_host failed")
        return b

    def write(r):
        if stage == "write" and r == 5:
            raise IOError("encode failed")

    t0 = time.time()
    with pytest.raises((OSError, RuntimeError, IOError)) as ei:
        pipeline.run_ordered_pipeline(read, process, write, depth=2)
    assert "failed" in str(ei.value)
    assert time.time() - t0 < 5.0
    time.sleep(0.2)
    assert not [t for t in threading.enumerate() if t.name in ("vex-reader", "vex-writer")]


def test_bad_depth():
    with pytest.raises(ValueError):
        pipeline.run_ordered_pipeline(lambda: None, lambda b, s: b, lambda r: None, depth=0)


def test_video_metadata_keys_and_fallbacks(tmp_path):
    """utils/video_utils.py:38-135 semantics: full key set, frame_count alias, zero record for an unreadable file."""
    cv2 = pytest.importorskip("cv2")
    from video_enhancer_proto_b200 import video_io
    assert video_io.parse_rate("30000/1001") == pytest.approx(29.97, abs=1e-2)
    assert video_io.parse_rate("25") == 25.0 and video_io.parse_rate("x/y") == 30.0 and video_io.parse_rate("1/0") == 30.0
    p = tmp_path / "c.mp4"
    wr = cv2.VideoWriter(str(p), cv2.VideoWriter_fourcc(*"mp4v"), 20.0, (64, 48))
    if not wr.isOpened():
        pytest.skip("OpenCV cannot write mp4v here")
    for i in range(7):
        wr.write(np.full((48, 64, 3), 30 * i, np.uint8))
    wr.release()
    m = video_io.video_metadata(str(p))
    assert set(m) == set(video_io.KEYS)
    assert (m["width"], m["height"]) == (64, 48) and m["frame_count"] == m["total_frames"] == 7
    assert m["fps"] == pytest.approx(20.0, abs=0.5) and m["duration"] == pytest.approx(0.35, abs=0.06) and m["file_size"] > 0
    z = video_io.video_metadata(str(tmp_path / "missing.mp4"))
    assert set(z) == set(video_io.KEYS) and z["frame_count"] == 0 and z["width"] == 0


@pytest.mark.parametrize("depth,inflight", [(1, 1), (2, 2), (3, 1)])
def test_submit_finish_split_keeps_batches_in_flight_and_buffers_safe(depth, inflight):
    """The asynchronous form the video loop uses (vex_submit / vex_wait): `process` submits and returns a handle,
    `finish` waits.  Up to `inflight` batches are submitted before the oldest is waited for; input buffers (filled by
    the reader thread) and output buffers (read by the writer thread) rotate over n_slots(depth, inflight) and must
    never be overwritten while a later stage still needs them, even with a slow 'GPU' and a slow writer."""
    slots = pipeline.n_slots(depth, inflight)
    ins = [np.zeros(4, np.int64) for _ in range(slots)]
    outs = [np.zeros(4, np.int64) for _ in range(slots)]
    n_batches, n_read = 45, 0
    events, got = [], []
    lock = threading.Lock()

    def read():
        nonlocal n_read
        if n_read == n_batches:
            return None
        ins[n_read % slots][:] = n_read                     # the reader writes straight into the rotating input buffer
        n_read += 1
        return n_read - 1

    class FakeGpu(threading.Thread):                        # completes submitted work asynchronously, in order
        def __init__(self):
            super().__init__(daemon=True)
            self.q, self.done = [], {}
            self.cv = threading.Condition()

        def run(self):
            while True:
                with self.cv:
                    while not self.q:
                        self.cv.wait()
                    k, slot = self.q.pop(0)
                if k is None:
                    return
                time.sleep(0.002)
                outs[slot][:] = ins[slot] * 10              # reads the input buffer LATE: it must still hold batch k
                with self.cv:
                    self.done[k] = True
                    self.cv.notify_all()

    gpu = FakeGpu()
    gpu.start()

    def submit(k, slot):
        assert slot == k % slots
        with lock:
            events.append(("submit", k))
        with gpu.cv:
            gpu.q.append((k, slot))
            gpu.cv.notify_all()
        return k, slot

    def finish(handle):
        k, slot = handle
        with gpu.cv:
            while k not in gpu.done:
                gpu.cv.wait()
        with lock:
            events.append(("finish", k))
        return k, outs[slot]

    def write(res):
        k, buf = res
        time.sleep(0.001)
        assert (buf == 10 * k).all(), (k, buf)
        got.append(k)

    n = pipeline.run_ordered_pipeline(read, submit, write, depth=depth, finish=finish, inflight=inflight)
    with gpu.cv:
        gpu.q.append((None, 0))
        gpu.cv.notify_all()
    assert n == n_batches and got == list(range(n_batches))
    # batch k + inflight was submitted before batch k was waited for (the GPU queue never drains between batches)
    pos = {e: i for i, e in enumerate(events)}
    for k in range(n_batches - inflight):
        assert pos[("submit", k + inflight)] < pos[("finish", k)], k


def test_submit_finish_error_still_waits_for_submitted_work():
    waited = []

    def submit(k, slot):
        if k == 3:
            raise RuntimeError("boom")
        return k

    def finish(k):
        waited.append(k)
        return k

    with pytest.raises(RuntimeError, match="boom"):
        pipeline.run_ordered_pipeline(make_reader(10), submit, lambda r: None, depth=2, finish=finish, inflight=2)
    assert sorted(waited) == [0, 1, 2]                      # everything handed to the GPU before the error was waited for
