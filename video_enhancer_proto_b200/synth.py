"""Synthetic frame sources (there are no real clips: the reference's demo mp4s are LFS stubs).

``demo_frame`` re-creates the reference's demo-clip recipe (app.py:1508-1554) in memory,
vectorised, without the lossy mp4v round trip; the other generators are the inputs
SURVEY.md section 8(c)/(d) lists for parity and benchmarking.
"""
from __future__ import annotations

import math
from typing import List

import numpy as np


def demo_frame(i: int, fps: int = 24, size=(480, 360), rng: np.random.Generator | None = None,
               bgr: bool = True) -> np.ndarray:
    """Frame ``i`` of the reference's demo video (app.py:1508-1554), uint8 HWC."""
    import cv2

    w, h = size
    t = i / fps
    x = np.arange(w, dtype=np.float64)[None, :]
    y = np.arange(h, dtype=np.float64)[:, None]
    # int() truncation of 127 + 127*sin(.) as in app.py:1518-1520
    r = np.trunc(127 + 127 * np.sin(x * 0.02 + t * 2)) + 0 * y
    g = np.trunc(127 + 127 * np.cos(y * 0.02 + t * 1.5)) + 0 * x
    b = np.trunc(127 + 127 * np.sin((x + y) * 0.01 + t))
    frame = np.stack([b, g, r], axis=-1).astype(np.uint8)           # BGR (app.py:1521)
    frame = np.ascontiguousarray(frame)

    cx = int((math.sin(t * 2.5) * 0.3 + 0.5) * w)
    cy = int((math.cos(t * 2) * 0.3 + 0.5) * h)
    cv2.circle(frame, (cx, cy), 30, (0, 255, 255), -1)
    rx = int((math.cos(t * 3) * 0.25 + 0.75) * (w - 80))
    ry = int((math.sin(t * 2.5) * 0.25 + 0.25) * (h - 60))
    cv2.rectangle(frame, (rx, ry), (rx + 80, ry + 60), (255, 0, 255), -1)
    ccx, ccy = w // 2, h // 2
    angle = t * 120
    ll = min(w, h) // 3
    ex = int(ccx + ll * math.cos(math.radians(angle)))
    ey = int(ccy + ll * math.sin(math.radians(angle)))
    cv2.line(frame, (ccx, ccy), (ex, ey), (0, 255, 0), 4)
    cv2.putText(frame, "ENHANCEMENT DEMO", (10, 30), cv2.FONT_HERSHEY_SIMPLEX, 0.8, (255, 255, 255), 2)
    cv2.putText(frame, f"Time: {t:.2f}s | Frame: {i + 1:03d}", (10, h - 40), cv2.FONT_HERSHEY_SIMPLEX, 0.6,
                (255, 255, 255), 2)
    cv2.putText(frame, f"Resolution: {w}x{h} | FPS: {fps}", (10, h - 10), cv2.FONT_HERSHEY_SIMPLEX, 0.5,
                (200, 200, 200), 1)
    if i % 8 == 0:                                                   # app.py:1551-1553
        rng = rng or np.random.default_rng(0)
        noise = rng.normal(0, 10, frame.shape).astype(np.int16)
        frame = np.clip(frame.astype(np.int16) + noise, 0, 255).astype(np.uint8)
    if not bgr:
        frame = np.ascontiguousarray(frame[:, :, ::-1])
    return frame


def random_frame(h: int, w: int, rng: np.random.Generator) -> np.ndarray:
    return rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)


def edges_frame(h: int, w: int, rng: np.random.Generator) -> np.ndarray:
    """Smooth gradients with hard 0/255 edges: bicubic overshoot exercises both clamps."""
    yy, xx = np.mgrid[0:h, 0:w]
    f = np.empty((h, w, 3), dtype=np.uint8)
    f[..., 0] = (xx * 255 // max(w - 1, 1)).astype(np.uint8)
    f[..., 1] = (yy * 255 // max(h - 1, 1)).astype(np.uint8)
    f[..., 2] = ((xx + yy) * 255 // max(w + h - 2, 1)).astype(np.uint8)
    for _ in range(12):
        y0, x0 = int(rng.integers(0, h)), int(rng.integers(0, w))
        hh, ww = int(rng.integers(1, max(2, h // 3))), int(rng.integers(1, max(2, w // 3)))
        f[y0:y0 + hh, x0:x0 + ww] = 255 if rng.integers(0, 2) else 0
    return f


def checker_frame(h: int, w: int) -> np.ndarray:
    yy, xx = np.mgrid[0:h, 0:w]
    c = (((xx + yy) & 1) * 255).astype(np.uint8)
    return np.repeat(c[..., None], 3, axis=2).copy()


def impulse_frame(h: int, w: int, y: int | None = None, x: int | None = None) -> np.ndarray:
    f = np.zeros((h, w, 3), dtype=np.uint8)
    f[h // 2 if y is None else y, w // 2 if x is None else x] = (255, 128, 64)
    return f


def constant_frame(h: int, w: int, v: int) -> np.ndarray:
    return np.full((h, w, 3), v, dtype=np.uint8)


def mixed_clip(n: int, h: int, w: int, seed: int = 1234) -> List[np.ndarray]:
    """BASELINE.md C2/C3 input: thirds of uniform-random, demo-pattern and gradient+edge frames."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        k = i % 3
        if k == 0:
            out.append(random_frame(h, w, rng))
        elif k == 1:
            out.append(demo_frame(i, size=(w, h), rng=rng, bgr=False))
        else:
            out.append(edges_frame(h, w, rng))
    return out


def tag_frame(frame: np.ndarray, index: int) -> np.ndarray:
    """Write a 32-bit frame index into the top-left 4x3 bytes... of the FIRST pixel rows.

    Used by the multi-GPU ordering check: the tag survives neither network, so the check is
    done on the INPUT side (frame i is known to go to rank r) and on the output order via
    ``read_tag`` of the echoed input index array kept by the host reorder queue.
    """
    f = frame.copy()
    f[0, 0, 0] = index & 0xFF
    f[0, 0, 1] = (index >> 8) & 0xFF
    f[0, 0, 2] = (index >> 16) & 0xFF
    f[0, 1, 0] = (index >> 24) & 0xFF
    return f


def read_tag(frame: np.ndarray) -> int:
    return int(frame[0, 0, 0]) | (int(frame[0, 0, 1]) << 8) | (int(frame[0, 0, 2]) << 16) | (int(frame[0, 1, 0]) << 24)
