"""Ordered three-stage frame pipeline: reader thread -> caller's thread (GPU) -> writer thread.

SURVEY.md §8(f)1: the reference's video loop (app.py:431-483, basic_upscaler_fallback.py:100-152) decodes a
frame, enhances it and encodes it on one thread, so the GPU idles during `cap.read()` and `out.write()`.  Here
decode and encode run on their own threads (OpenCV releases the GIL inside both) while the calling thread keeps the
GPU busy; batches are written in the order they were read.  Nothing in this module touches CUDA.
"""
from __future__ import annotations

import collections
import queue
import threading
from typing import Callable, Optional, TypeVar

T = TypeVar("T")
R = TypeVar("R")

_END = object()


def n_slots(depth: int, inflight: int = 0) -> int:
    """Buffers a stage must rotate over: `depth` batches queued, one being written (or read) and one being produced
    (or consumed), plus `inflight` batches submitted to the GPU but not yet waited for, can all be alive at once.
    The same count holds on the input side (reader -> queue -> submit -> finish) and on the output side
    (submit -> finish -> queue -> writer)."""
    return depth + 2 + inflight


def run_ordered_pipeline(read_batch: Callable[[], Optional[T]], process: Callable[[T, int], R],
                         write: Callable[[R], None], depth: int = 2,
                         finish: Optional[Callable] = None, inflight: int = 1) -> int:
    """Runs `read_batch` (returns None at end of stream) on a reader thread, `process(batch, slot)` on the calling
    thread and `write(result)` on a writer thread.  `slot` cycles over ``n_slots(depth)`` values; a result may alias
    a per-slot buffer because the slot is not handed out again before the writer is done with it.  Returns the number
    of batches processed.  The first exception raised by any stage stops all three and is re-raised here.

    With `finish`, `process` is an asynchronous SUBMIT that returns a handle (vex_submit's ticket) and
    `finish(handle)` blocks for the result (vex_wait): up to `inflight` submitted batches are kept un-waited, so batch
    k+1 is submitted (its H2D and kernels queued) before batch k is waited for and the GPU never drains between
    batches.  `slot` then cycles over ``n_slots(depth, inflight)`` values."""
    if depth < 1:
        raise ValueError("depth must be >= 1")
    if finish is not None and inflight < 1:
        raise ValueError("inflight must be >= 1")
    q_in: "queue.Queue" = queue.Queue(maxsize=depth)
    q_out: "queue.Queue" = queue.Queue(maxsize=depth)
    stop = threading.Event()
    errors: list = []

    def put(q, item) -> bool:
        while not stop.is_set():
            try:
                q.put(item, timeout=0.05)
                return True
            except queue.Full:
                continue
        return False

    def get(q):
        while not stop.is_set():
            try:
                return q.get(timeout=0.05)
            except queue.Empty:
                continue
        return _END

    def reader():
        try:
            while not stop.is_set():
                b = read_batch()
                if b is None:
                    break
                if not put(q_in, b):
                    return
        except BaseException as e:          # noqa: BLE001 - re-raised on the calling thread
            errors.append(e)
            stop.set()
        finally:
            put(q_in, _END)

    def writer():
        try:
            while True:
                r = get(q_out)
                if r is _END:
                    return
                write(r)
        except BaseException as e:          # noqa: BLE001
            errors.append(e)
            stop.set()

    tr = threading.Thread(target=reader, name="vex-reader", daemon=True)
    tw = threading.Thread(target=writer, name="vex-writer", daemon=True)
    tr.start()
    tw.start()
    done = 0
    slots = n_slots(depth, inflight if finish is not None else 0)
    pending: "collections.deque" = collections.deque()
    try:
        while True:
            b = get(q_in)
            if b is _END:
                break
            r = process(b, done % slots)
            done += 1
            if finish is not None:
                pending.append(r)
                if len(pending) <= inflight:
                    continue
                r = finish(pending.popleft())
            if not put(q_out, r):
                break
        while pending and not stop.is_set():
            if not put(q_out, finish(pending.popleft())):
                break
    except BaseException as e:              # noqa: BLE001
        errors.append(e)
        stop.set()
    finally:
        while pending:                      # after an error: still wait for what the GPU was handed (its buffers are reused)
            try:
                finish(pending.popleft())
            except BaseException:           # noqa: BLE001
                pass
        put(q_out, _END)
        tw.join()
        stop.set()                          # releases a reader blocked on a full queue after an early exit
        tr.join()
    if errors:
        raise errors[0]
    return done
