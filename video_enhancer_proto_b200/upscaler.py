"""Host-side mirror of the reference's upscaler/enhancer interface over libvex.so.

One class, ``B200VideoEnhancer``, answers both call shapes the reference uses for this path:

* ``GPUVideoEnhancer`` (app.py:242-621): ``load_models``, ``models_loaded``, ``enhance_frame_gpu``,
  ``_enhance_frame_with_gpu``, ``_enhance_frame_cpu_fallback``, ``_process_frame_batch_gpu``,
  ``process_video`` -> ``(ok, message)``;
* the handler protocol the agents dispatch on (``models/enhancement/frame/basic_upscaler_fallback.py``
  :45,67,171; ``agents/enhancer/video_enhancer_sota.py:316-366``): ``__init__(device, model_name,
  scale)``, ``enhance_video`` / ``restore_video`` -> stats dict, ``get_model_info``.

All pixel arithmetic runs in hand-written sm_100a kernels behind the C ABI (include/vex.h).
PyTorch is used only for device/pinned buffers and stream handles.  There is no CPU fallback:
``_enhance_frame_cpu_fallback`` keeps its reference name for drop-in compatibility but runs the
bicubic kernel on the GPU, and every method raises (or, for ``process_video``, returns
``(False, msg)`` like app.py:514-519) if the library or device is unavailable.
"""
from __future__ import annotations

import ctypes as C
import logging
import time
from pathlib import Path
from typing import Dict, List, Mapping, Optional, Sequence

import numpy as np
import torch

from . import _lib, pipeline, video_io, weights as W

logger = logging.getLogger(__name__)


class B200VideoEnhancer:
    """Drop-in for ``GPUVideoEnhancer`` / a frame handler.  NOT thread-safe; one instance per GPU."""

    def __init__(self, device: str = "cuda", model_name: str = "AdvancedUpscaler", scale: int = 2,
                 state_dict: Optional[Mapping[str, object]] = None, seed: Optional[int] = None,
                 batch_size: int = 8):
        if scale != 2:
            raise ValueError("B200VideoEnhancer implements the reference's fixed 2x path (app.py:280)")
        self.device = device
        self.model_name = model_name
        self.scale = scale
        self.batch_size = batch_size                      # app.py:453
        self.pipeline_depth = 2                           # batches queued either side of the GPU in the video loop
        self.models_loaded = False
        self.processing_stats = {"total_processed": 0, "total_time": 0.0}   # app.py:236-240
        self.job_history: List[Dict] = []
        self._ctx = C.c_void_p()
        self._lib = None
        self._state = None
        self._pending_state = (state_dict, seed)
        self._widths, self._up_ch, self._n_extra = W.REFERENCE_WIDTHS, W.REFERENCE_UP_CH, 0
        self._pinned = {}                                  # (tag, nbytes) -> pinned tensors
        self._dev = {}

    # ------------------------------------------------------------------ lifetime
    def _ordinal(self) -> int:
        d = torch.device(self.device if self.device != "cuda" else "cuda:0")
        if d.type != "cuda":
            raise RuntimeError(f"B200VideoEnhancer needs a CUDA device, got {self.device!r} (no CPU fallback)")
        return d.index or 0

    def _ensure_ctx(self):
        if self._ctx:
            return
        self._lib = _lib.load()
        rc = self._lib.vex_create(self._ordinal(), C.byref(self._ctx))
        if rc != 0:
            msg = self._lib.vex_last_error(None)
            self._ctx = C.c_void_p()
            raise _lib.VexError(f"vex_create failed ({rc}): {msg.decode() if msg else '?'}")

    def load_models(self) -> bool:
        """app.py:252-263.  Builds the context and uploads the CNN weights."""
        try:
            self._ensure_ctx()
            self._ensure_weights()
            self.models_loaded = True
            return True
        except Exception as e:  # reference returns False and logs
            logger.error(f"Model configuration failed: {e}")
            return False

    def load_state_dict(self, state_dict: Mapping[str, object]) -> None:
        """Upload a checkpoint with the reference module's key names."""
        self._ensure_ctx()
        widths, up_ch, n_extra = W.infer_plan(state_dict)
        st = W.normalize_state(state_dict, widths, up_ch, n_extra)
        desc = _lib.VexCnnDesc(widths[0], widths[1], widths[2], widths[3], up_ch, 1e-5, n_extra)
        ptrs = (C.c_void_p * len(st))(*[a.ctypes.data for a in st.values()])
        rc = self._lib.vex_cnn_load_weights(self._ctx, C.byref(desc), ptrs, len(st))
        _lib.check(self._ctx, rc, "vex_cnn_load_weights")
        self._state, self._widths, self._up_ch, self._n_extra = st, widths, up_ch, n_extra

    def state_dict(self):
        self._ensure_weights()
        return self._state

    def _ensure_weights(self):
        if self._state is not None:
            return
        sd, seed = self._pending_state
        self.load_state_dict(sd if sd is not None else W.init_state(seed))   # app.py:303-311

    def close(self):
        if self._ctx:
            self._lib.vex_destroy(self._ctx)
            self._ctx = C.c_void_p()
        self._pinned.clear()
        self._dev.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ buffers
    def _device_buf(self, tag: str, nbytes: int) -> torch.Tensor:
        t = self._dev.get(tag)
        if t is None or t.numel() < nbytes:
            t = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{self._ordinal()}")
            self._dev[tag] = t
        return t

    def _pinned_buf(self, tag: str, nbytes: int) -> torch.Tensor:
        t = self._pinned.get(tag)
        if t is None or t.numel() < nbytes:
            t = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
            self._pinned[tag] = t
        return t

    @staticmethod
    def _check_frames(frames: Sequence[np.ndarray]):
        if len(frames) == 0:
            raise ValueError("empty frame batch")
        h, w = frames[0].shape[:2]
        for f in frames:
            if f.dtype != np.uint8 or f.ndim != 3 or f.shape[2] != 3 or f.shape[:2] != (h, w):
                raise ValueError("frames must be uint8 HWC with 3 channels and one common size")
        return h, w

    def _run_host(self, mode: int, frames: Sequence[np.ndarray], flags: int, slot: int = 0) -> np.ndarray:
        """frames (host) -> [n, 2h, 2w, 3] uint8 (host) through vex_upscale_host.  The result is a view of the
        pinned output buffer of `slot`; it stays valid until the next call with the same slot."""
        self._ensure_ctx()
        if mode == _lib.VEX_MODE_CNN:
            self._ensure_weights()
        h, w = self._check_frames(frames)
        n = len(frames)
        fin = h * w * 3
        pin_in = self._pinned_buf("in", n * fin)
        pin_out = self._pinned_buf(f"out{slot}", n * fin * 4)
        vin = pin_in.numpy()[: n * fin].reshape(n, h, w, 3)
        for i, f in enumerate(frames):
            vin[i] = f
        rc = self._lib.vex_upscale_host(self._ctx, mode, pin_in.data_ptr(), pin_out.data_ptr(), n, h, w,
                                        min(n, self.batch_size), flags)
        _lib.check(self._ctx, rc, "vex_upscale_host")
        return pin_out.numpy()[: n * fin * 4].reshape(n, 2 * h, 2 * w, 3)

    # ------------------------------------------------------------------ device-resident entry points
    def upscale_device(self, frames_u8: torch.Tensor, mode: str = "cnn", in_bgr: bool = False,
                       out_bgr: bool = False, round_mode: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[n,h,w,3] uint8 CUDA tensor -> [n,2h,2w,3] uint8 CUDA tensor, asynchronous on the current stream."""
        self._ensure_ctx()
        if frames_u8.dtype != torch.uint8 or frames_u8.dim() != 4 or frames_u8.shape[3] != 3 or not frames_u8.is_cuda:
            raise ValueError("expected a CUDA uint8 tensor [n,h,w,3]")
        frames_u8 = frames_u8.contiguous()
        n, h, w, _ = frames_u8.shape
        if out is None:
            out = torch.empty((n, 2 * h, 2 * w, 3), dtype=torch.uint8, device=frames_u8.device)
        stream = torch.cuda.current_stream(frames_u8.device).cuda_stream
        if mode == "cnn":
            self._ensure_weights()
            rc = self._lib.vex_cnn2x_u8(self._ctx, frames_u8.data_ptr(), out.data_ptr(), n, h, w,
                                        int(in_bgr), int(out_bgr), stream)
            _lib.check(self._ctx, rc, "vex_cnn2x_u8")
        elif mode == "bicubic":
            rc = self._lib.vex_bicubic2x_u8(self._ctx, frames_u8.data_ptr(), out.data_ptr(), n, h, w,
                                            int(round_mode), stream)
            _lib.check(self._ctx, rc, "vex_bicubic2x_u8")
        else:
            raise ValueError(f"unknown mode {mode!r}")
        return out

    def cnn_debug_f32(self, frames_u8: torch.Tensor, tap: int = 0, in_bgr: bool = False) -> torch.Tensor:
        """Parity hook (vex_cnn2x_f32_debug): fp32 NCHW of the selected activation."""
        self._ensure_ctx()
        self._ensure_weights()
        frames_u8 = frames_u8.contiguous()
        n, h, w, _ = frames_u8.shape
        c1, c2, c3, c4 = self._widths
        shape = {0: (n, 3, 2 * h, 2 * w), 1: (n, c1, h, w), 2: (n, c2, h, w), 3: (n, c3, h, w),
                 4: (n, c4, h, w), 5: (n, 3, h, w), 6: (n, self._up_ch, 2 * h, 2 * w)}.get(tap, (n, c3, h, w))
        if tap > 6 + self._n_extra:
            raise ValueError(f"tap {tap}: the plan has {self._n_extra} extra block(s)")
        out = torch.empty(shape, dtype=torch.float32, device=frames_u8.device)
        rc = self._lib.vex_cnn2x_f32_debug(self._ctx, frames_u8.data_ptr(), out.data_ptr(), n, h, w,
                                           int(in_bgr), tap, torch.cuda.current_stream(frames_u8.device).cuda_stream)
        _lib.check(self._ctx, rc, "vex_cnn2x_f32_debug")
        return out

    LAYER_NAMES = ("conv1", "conv2", "conv3", "conv4", "conv5", "upconv1", "upconv2")

    def cnn_profile_ms(self) -> Dict[str, float]:
        """Per-layer device ms of the last CNN call made with ``set_option("profile", 1)`` (vex_cnn_profile_ms)."""
        self._ensure_ctx()
        ms = (C.c_float * 7)()
        rc = self._lib.vex_cnn_profile_ms(self._ctx, ms)
        _lib.check(self._ctx, rc, "vex_cnn_profile_ms")
        return dict(zip(self.LAYER_NAMES, [float(v) for v in ms]))

    def kernel_launches(self) -> int:
        self._ensure_ctx()
        v = C.c_ulonglong()
        self._lib.vex_kernel_launches(self._ctx, C.byref(v))
        return int(v.value)

    def set_option(self, name: str, value: int) -> int:
        self._ensure_ctx()
        return self._lib.vex_set_option(self._ctx, name.encode(), int(value))

    # ------------------------------------------------------------------ GPUVideoEnhancer surface
    def enhance_frame_gpu(self, frame):
        """app.py:314-319: uint8 HWC ndarray (or fp32 NCHW tensor in [0,1]) -> uint8 HWC (2H,2W,3)."""
        return self._enhance_frame_with_gpu(frame)

    def _enhance_frame_with_gpu(self, frame):
        """app.py:322-352: single frame through the CNN, no colour swap."""
        if isinstance(frame, torch.Tensor):               # app.py:337-338: fp32 NCHW in [0,1]
            t = frame.detach()
            if t.dim() == 4:
                t = t[0]
            frame = (t.permute(1, 2, 0).clamp(0, 1) * 255.0).round().to(torch.uint8).cpu().numpy()
        return self._run_host(_lib.VEX_MODE_CNN, [np.ascontiguousarray(frame)], 0)[0].copy()

    def _enhance_frame_cpu_fallback(self, frame):
        """app.py:359-385 semantics (bicubic, clip, truncate) — computed by the GPU kernel."""
        if isinstance(frame, torch.Tensor):
            t = frame.detach()
            if t.dim() == 4:
                t = t[0]
            frame = (t.permute(1, 2, 0).clamp(0, 1) * 255.0).round().to(torch.uint8).cpu().numpy()
        return self._run_host(_lib.VEX_MODE_BICUBIC, [np.ascontiguousarray(frame)], 0)[0].copy()

    def _process_frame_batch_gpu(self, frames_batch: List[np.ndarray]) -> List[np.ndarray]:
        """app.py:529-553: RGB uint8 frames in, BGR uint8 (2H,2W,3) frames out."""
        out = self._run_host(_lib.VEX_MODE_CNN, frames_batch, flags=0b10)
        return [out[i].copy() for i in range(out.shape[0])]

    def process_video(self, input_path, output_path, target_fps=30):
        """app.py:388-519: never raises; returns (ok, message)."""
        try:
            start = time.time()
            stats = self._run_video(str(input_path), str(output_path), target_fps=target_fps)
            dt = time.time() - start
            self.processing_stats["total_processed"] += 1          # app.py:509-510
            self.processing_stats["total_time"] += dt
            self.job_history.append({"engine": "B200 Upscaler", "model": self.model_name,
                                     "frames": stats["frames_processed"], "time": f"{dt:.1f}s",
                                     "input": str(input_path), "output": str(output_path)})
            return True, f"✅ Enhanced {stats['frames_processed']} frames in {dt:.1f}s with B200 acceleration"
        except Exception as e:
            logger.error(f"B200 video processing failed: {e}")
            return False, f"❌ B200 processing failed: {str(e)}"

    # ------------------------------------------------------------------ handler protocol
    def enhance_video(self, input_path: str, output_path: str, **kwargs) -> Dict:
        """basic_upscaler_fallback.py:67-169 contract: raises ValueError on unreadable input."""
        return self._run_video(str(input_path), str(output_path), target_fps=kwargs.get("target_fps"))

    def restore_video(self, input_path: str, output_path: str, **kwargs) -> Dict:
        """vsr/realesrgan_handler.py:214-273 key set."""
        s = self.enhance_video(input_path, output_path, **kwargs)
        s.update(processing_mode="frame_batch", scale_factor=self.scale)
        return s

    def get_model_info(self) -> Dict:
        c1, c2, c3, c4 = self._widths
        return {
            "name": "B200Upscaler",
            "description": "AdvancedUpscaler 2x CNN (app.py:269-301) / bicubic 2x on hand-written sm_100a kernels",
            "scale": self.scale,
            "device": self.device,
            "model": self.model_name,
            "architecture": f"conv3x3 3-{c1}-{c2}(+skip)-{c3}" + f"-{c3}" * self._n_extra +
                            f"-{c4}-3, ConvT k4s2 3-{self._up_ch}, ConvT k3 {self._up_ch}-3",
            "parameters": int(sum(int(np.prod(s)) for k, s in
                                  W.expected_shapes(self._widths, self._up_ch, self._n_extra).items()
                                  if "running" not in k)),
            "capabilities": ["upscaling", "bicubic_fallback"],
            "precision": "bf16 operands, fp32 accumulate",
        }

    # ------------------------------------------------------------------ video loop (app.py:431-483)
    def _use_bicubic(self) -> bool:
        return self.model_name.upper() in ("CUBIC", "BICUBIC")

    def _run_video(self, input_path: str, output_path: str, target_fps=None) -> Dict:
        import cv2

        cap = cv2.VideoCapture(input_path)
        if not cap.isOpened():
            raise ValueError(f"Could not open video: {input_path}")
        out = None
        try:
            fps = cap.get(cv2.CAP_PROP_FPS)
            width = int(cap.get(cv2.CAP_PROP_FRAME_WIDTH))
            height = int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT))
            Path(output_path).parent.mkdir(parents=True, exist_ok=True)
            fourcc = cv2.VideoWriter_fourcc(*"mp4v")
            out_fps = min(target_fps, fps) if target_fps else fps            # app.py:447
            out = cv2.VideoWriter(output_path, fourcc, out_fps, (width * 2, height * 2))
            if not out.isOpened():
                raise ValueError(f"Could not create output video: {output_path}")
            mode = _lib.VEX_MODE_BICUBIC if self._use_bicubic() else _lib.VEX_MODE_CNN
            # cv2 frames are BGR.  bicubic is channel-agnostic (basic_upscaler_fallback.py:142); for the CNN the
            # frames stay BGR on the host and in_bgr|out_bgr do app.py:469 and :550 inside the kernels
            flags = 1 if mode == _lib.VEX_MODE_BICUBIC else 0b11
            n_done = 0

            def read_batch():
                batch: List[np.ndarray] = []
                while len(batch) < self.batch_size:
                    ret, frame = cap.read()
                    if not ret:
                        break
                    batch.append(frame)
                return batch or None

            def process(batch, slot):
                return self._run_host(mode, batch, flags, slot=slot)

            def write(res):
                nonlocal n_done
                for i in range(res.shape[0]):
                    out.write(res[i])
                n_done += res.shape[0]

            # decode and encode on their own threads, the GPU call on this one
            pipeline.run_ordered_pipeline(read_batch, process, write, depth=self.pipeline_depth)
        finally:
            cap.release()
            if out is not None:
                out.release()
        return {
            "frames_processed": n_done, "frames_failed": 0, "input_frames": n_done, "output_frames": n_done,
            "scale": self.scale, "model": f"B200Upscaler-{self.model_name}",
            "input_resolution": f"{width}x{height}", "output_resolution": f"{width * 2}x{height * 2}",
            "fps": float(out_fps),
            "input_metadata": video_io.video_metadata(input_path),       # utils/video_utils.py:38 key set
        }
