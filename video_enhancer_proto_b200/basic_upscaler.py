"""GPU twin of ``BasicUpscalerFallback`` (models/enhancement/frame/basic_upscaler_fallback.py:42-181), SURVEY.md §8(f)3.

Same constructor order, ``enhance_video`` contract and ``get_model_info`` keys; the per-frame OpenCV calls
(basic_upscaler_fallback.py:113-128) run as CUDA kernels behind the C ABI:

    LANCZOS:  cv2.bilateralFilter(frame, 5, 50, 50) -> cv2.resize(INTER_LANCZOS4) -> cv2.filter2D(0.1 * sharpen) -> clip
    CUBIC:    cv2.resize(INTER_CUBIC)                                             -> cv2.filter2D(0.1 * sharpen) -> clip

The reference's sharpening kernel sums to 0.1, not 1 (``kernel * 0.1`` scales the centre tap too), so its output is the
sharpened image at a tenth of the brightness; that is what the reference writes and what is reproduced here.
    LINEAR / AREA / NEAREST:  cv2.resize(INTER_*) = vex_resize2x_u8 (bit-exact)       -> cv2.filter2D(0.1 * sharpen) -> clip
"""
from __future__ import annotations

import ctypes as C
import logging
from pathlib import Path
from typing import Dict, List, Sequence

import numpy as np
import torch

from . import _lib, pipeline
from .upscaler import B200VideoEnhancer

logger = logging.getLogger(__name__)

# basic_upscaler_fallback.py:123-124, computed in float32 exactly like numpy does there
SHARPEN_KERNEL = (np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]], dtype=np.float32) * 0.1).astype(np.float32)
BILATERAL_ARGS = (5, 50.0, 50.0)            # basic_upscaler_fallback.py:117


class B200BasicUpscaler:
    SUPPORTED = ("LANCZOS", "CUBIC", "LINEAR", "AREA", "NEAREST")          # basic_upscaler_fallback.py:51-57
    CV_INTERPOLATION = {"NEAREST": 0, "LINEAR": 1, "CUBIC": 2, "AREA": 3, "LANCZOS": 4}   # cv2.INTER_* values

    def __init__(self, device: str = "cuda", model_name: str = "LANCZOS", scale: int = 2, batch_size: int = 8):
        if scale != 2:
            raise ValueError("B200BasicUpscaler implements the 2x ratio the application uses")
        name = model_name.upper()
        # unknown names fall back to Lanczos like the reference's dict.get default (basic_upscaler_fallback.py:59-62)
        self.method = name if name in self.SUPPORTED else "LANCZOS"
        # the bilateral pre-filter is gated on the NAME, not on the interpolation that was picked (:115-117): an unknown
        # name gets the Lanczos resize without denoising
        self._denoise = name == "LANCZOS"
        self.device = device
        self.model_name = model_name
        self.scale = scale
        self.batch_size = batch_size
        self.pipeline_depth = 2
        self._enh = B200VideoEnhancer(device=device, model_name="bicubic", batch_size=batch_size)   # ctx + buffers
        self._k9 = (C.c_float * 9)(*[float(v) for v in SHARPEN_KERNEL.ravel()])

    def close(self):
        self._enh.close()

    # ------------------------------------------------------------------ device path
    def upscale_device(self, frames_u8: torch.Tensor, sharpen: bool = True) -> torch.Tensor:
        """[n,h,w,3] uint8 CUDA tensor (any channel order) -> [n,2h,2w,3] uint8: the reference's per-frame pipeline."""
        e = self._enh
        e._ensure_ctx()
        lib, ctx = e._lib, e._ctx
        x = frames_u8.contiguous()
        n, h, w, _ = x.shape
        st = torch.cuda.current_stream(x.device).cuda_stream
        if self.method == "LANCZOS":
            den = x
            if self._denoise:
                den = torch.empty_like(x)
                _lib.check(ctx, lib.vex_bilateral5_u8(ctx, x.data_ptr(), den.data_ptr(), n, h, w, BILATERAL_ARGS[1],
                                                      BILATERAL_ARGS[2], st), "vex_bilateral5_u8")
            up = torch.empty((n, 2 * h, 2 * w, 3), dtype=torch.uint8, device=x.device)
            _lib.check(ctx, lib.vex_lanczos4_2x_u8(ctx, den.data_ptr(), up.data_ptr(), n, h, w, st), "vex_lanczos4_2x_u8")
        elif self.method == "CUBIC":
            up = e.upscale_device(x, mode="bicubic", round_mode=1)          # cv2.resize INTER_CUBIC rounds to nearest
        else:                                                               # LINEAR / AREA / NEAREST (:54-56)
            up = torch.empty((n, 2 * h, 2 * w, 3), dtype=torch.uint8, device=x.device)
            _lib.check(ctx, lib.vex_resize2x_u8(ctx, x.data_ptr(), up.data_ptr(), n, h, w,
                                                self.CV_INTERPOLATION[self.method], st), "vex_resize2x_u8")
        if not sharpen:
            return up
        out = torch.empty_like(up)
        _lib.check(ctx, lib.vex_filter3x3_u8(ctx, up.data_ptr(), out.data_ptr(), n, 2 * h, 2 * w, self._k9, st),
                   "vex_filter3x3_u8")
        return out

    def enhance_frames(self, frames: Sequence[np.ndarray], slot: int = 0) -> List[np.ndarray]:
        """Host frames through the pipeline: pinned staging on both sides (asynchronous H2D / D2H on the current stream,
        one synchronisation at the end).  The returned arrays are views of the pinned result buffer of `slot`: valid until
        the next call with that slot (the video loop rotates slots, other callers copy what they keep)."""
        h, w = B200VideoEnhancer._check_frames(frames)
        n, fin = len(frames), h * w * 3
        e = self._enh
        pin_in = e._pinned_buf("b_in", n * fin)
        pin_out = e._pinned_buf(f"b_out{slot}", n * fin * 4)
        vin = pin_in.numpy()[: n * fin].reshape(n, h, w, 3)
        for i, f in enumerate(frames):
            vin[i] = f
        dev = torch.device(e.device if e.device != "cuda" else "cuda:0")
        x = e._device_buf("b_in", n * fin)[: n * fin].view(n, h, w, 3)
        x.copy_(pin_in[: n * fin].view(n, h, w, 3), non_blocking=True)
        y = self.upscale_device(x)
        pin_out[: n * fin * 4].view(n, 2 * h, 2 * w, 3).copy_(y, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        res = pin_out.numpy()[: n * fin * 4].reshape(n, 2 * h, 2 * w, 3)
        return [res[i] for i in range(n)]

    # ------------------------------------------------------------------ handler protocol
    def enhance_video(self, input_path: str, output_path: str) -> Dict:
        """basic_upscaler_fallback.py:67-169: same errors, side effects and statistics keys."""
        import cv2

        cap = cv2.VideoCapture(str(input_path))
        if not cap.isOpened():
            raise ValueError(f"Could not open video: {input_path}")                 # :84-85
        out = None
        processed = 0
        try:
            fps = cap.get(cv2.CAP_PROP_FPS)
            width = int(cap.get(cv2.CAP_PROP_FRAME_WIDTH))
            height = int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT))
            out_w, out_h = width * self.scale, height * self.scale
            Path(output_path).parent.mkdir(parents=True, exist_ok=True)              # :98
            out = cv2.VideoWriter(str(output_path), cv2.VideoWriter_fourcc(*"mp4v"), fps, (out_w, out_h))

            def read_batch():
                batch = []
                while len(batch) < self.batch_size:
                    ret, frame = cap.read()
                    if not ret:
                        break
                    batch.append(frame)
                return batch or None

            def process(batch, slot):
                return self.enhance_frames(batch, slot=slot)

            def write(res):
                nonlocal processed
                for f in res:
                    out.write(f)
                processed += len(res)

            pipeline.run_ordered_pipeline(read_batch, process, write, depth=self.pipeline_depth)
        finally:
            cap.release()
            if out is not None:
                out.release()
        return {
            "frames_processed": processed, "frames_failed": 0, "scale": self.scale,
            "model": f"BasicUpscaler-{self.model_name}",
            "input_resolution": f"{width}x{height}", "output_resolution": f"{out_w}x{out_h}",
            "interpolation_method": self.model_name,
        }

    def get_model_info(self) -> Dict:
        return {
            "name": "BasicUpscaler",
            "description": f"B200 CUDA upscaling using {self.model_name} (OpenCV-compatible arithmetic)",
            "scale": self.scale,
            "device": self.device,
            "interpolation_method": self.model_name,
            "capabilities": ["upscaling", "basic_enhancement"],
        }
