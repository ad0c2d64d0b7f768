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

    def cnn_profile_stats(self) -> Dict[str, Dict[str, float]]:
        """Mean / min / max per-layer device ms over every forward since ``set_option("profile", 1)``
        (vex_cnn_profile_stats) plus ``n`` = forwards counted."""
        self._ensure_ctx()
        mean, lo, hi, n = (C.c_float * 7)(), (C.c_float * 7)(), (C.c_float * 7)(), C.c_int()
        rc = self._lib.vex_cnn_profile_stats(self._ctx, mean, lo, hi, C.byref(n))
        _lib.check(self._ctx, rc, "vex_cnn_profile_stats")
        return {"mean": dict(zip(self.LAYER_NAMES, map(float, mean))), "min": dict(zip(self.LAYER_NAMES, map(float, lo))),
                "max": dict(zip(self.LAYER_NAMES, map(float, hi))), "n": int(n.value)}

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

    def _tensor_frame(self, frame: torch.Tensor) -> torch.Tensor:
        """app.py:337-338 / :367-368: the tensor branch takes fp32 NCHW (or CHW) in [0,1] as it is - no rounding to
        uint8 on the way in (the reference feeds it to conv1 / F.interpolate unquantised)."""
        t = frame.detach()
        if t.dim() == 3:
            t = t.unsqueeze(0)
        if t.dim() != 4 or t.shape[1] != 3:
            raise ValueError("tensor frames must be [1,3,H,W] or [3,H,W]")
        return t.to(device=f"cuda:{self._ordinal()}", dtype=torch.float32).contiguous()

    def _enhance_frame_with_gpu(self, frame):
        """app.py:322-352: single frame through the CNN, no colour swap."""
        if isinstance(frame, torch.Tensor):
            self._ensure_ctx()
            self._ensure_weights()
            t = self._tensor_frame(frame)
            n, _, h, w = t.shape
            out = torch.empty((n, 2 * h, 2 * w, 3), dtype=torch.uint8, device=t.device)
            rc = self._lib.vex_cnn2x_f32in(self._ctx, t.data_ptr(), out.data_ptr(), None, n, h, w, 0,
                                           torch.cuda.current_stream(t.device).cuda_stream)
            _lib.check(self._ctx, rc, "vex_cnn2x_f32in")
            return out[0].cpu().numpy()                    # app.py:345: squeeze(0) -> HWC
        return self._run_host(_lib.VEX_MODE_CNN, [np.ascontiguousarray(frame)], 0)[0].copy()

    def cnn_f32in_debug(self, frames_nchw: torch.Tensor) -> torch.Tensor:
        """Parity hook for the tensor branch: fp32 NCHW in -> pre-quantisation fp32 NCHW [n,3,2h,2w] (vex_cnn2x_f32in)."""
        self._ensure_ctx()
        self._ensure_weights()
        t = self._tensor_frame(frames_nchw)
        n, _, h, w = t.shape
        out = torch.empty((n, 3, 2 * h, 2 * w), dtype=torch.float32, device=t.device)
        rc = self._lib.vex_cnn2x_f32in(self._ctx, t.data_ptr(), None, out.data_ptr(), n, h, w, 0,
                                       torch.cuda.current_stream(t.device).cuda_stream)
        _lib.check(self._ctx, rc, "vex_cnn2x_f32in")
        return out

    def _enhance_frame_cpu_fallback(self, frame):
        """app.py:359-385 semantics (bicubic, clip, truncate) — computed by the GPU kernel."""
        if isinstance(frame, torch.Tensor):
            self._ensure_ctx()
            t = self._tensor_frame(frame)
            n, _, h, w = t.shape
            out = torch.empty((n, 2 * h, 2 * w, 3), dtype=torch.uint8, device=t.device)
            rc = self._lib.vex_bicubic2x_f32in(self._ctx, t.data_ptr(), out.data_ptr(), n, h, w,
                                               torch.cuda.current_stream(t.device).cuda_stream)
            _lib.check(self._ctx, rc, "vex_bicubic2x_f32in")
            return out[0].cpu().numpy()
        return self._run_host(_lib.VEX_MODE_BICUBIC, [np.ascontiguousarray(frame)], 0)[0].copy()

    def _process_frame_batch_gpu(self, frames_batch: List[np.ndarray]) -> List[np.ndarray]:
        """app.py:529-553: RGB uint8 frames in, BGR uint8 (2H,2W,3) frames out.

        The frames of the list are separate pageable arrays; ``vex_upscale_frames`` stages them chunk by chunk into
        the ctx's pinned buffers while the previous chunk computes, and the results land in a FRESH pinned tensor
        (torch's caching host allocator) whose rows are returned as the output arrays - no host copy on the way out,
        and the arrays stay valid for as long as the caller keeps them."""
        return self._run_frames(_lib.VEX_MODE_CNN, frames_batch, flags=0b10)

    def _run_frames(self, mode: int, frames: Sequence[np.ndarray], flags: int) -> List[np.ndarray]:
        self._ensure_ctx()
        if mode == _lib.VEX_MODE_CNN:
            self._ensure_weights()
        h, w = self._check_frames(frames)
        n = len(frames)
        keep = [f if f.flags["C_CONTIGUOUS"] else np.ascontiguousarray(f) for f in frames]
        ptrs = (C.c_void_p * n)(*[f.ctypes.data for f in keep])
        out = torch.empty((n, 2 * h, 2 * w, 3), dtype=torch.uint8, pin_memory=True)
        rc = self._lib.vex_upscale_frames(self._ctx, mode, ptrs, out.data_ptr(), n, h, w, self.batch_size, flags)
        _lib.check(self._ctx, rc, "vex_upscale_frames")
        res = out.numpy()
        return [res[i] for i in range(n)]

    def process_video(self, input_path, output_path, target_fps=30):
        """app.py:388-519: never raises; returns (ok, message)."""
        try:
            start = time.time()
            stats = self._run_video(str(input_path), str(output_path), target_fps=target_fps)
            stats.pop("_size", None)
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
        s = self._run_video(str(input_path), str(output_path), target_fps=kwargs.get("target_fps"))
        if not kwargs.get("_keep_size"):
            s.pop("_size", None)
        return s

    def restore_video(self, input_path: str, output_path: str, **kwargs) -> Dict:
        """vsr/realesrgan_handler.py:214-273 key set."""
        s = self.enhance_video(input_path, output_path, _keep_size=True, **kwargs)
        w, h = s.pop("_size")
        # realesrgan_handler.py:265-273: resolutions are (H, W) tuples here (the frame handlers' are "WxH" strings)
        s.update(processing_mode="b200_frame_batch", scale_factor=self.scale, input_resolution=(h, w),
                 output_resolution=(h * self.scale, w * self.scale))
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
            if width <= 0 or height <= 0:
                raise ValueError(f"Could not read the frame size of: {input_path}")
            n_done = 0
            self._ensure_ctx()
            if mode == _lib.VEX_MODE_CNN:
                self._ensure_weights()
            # Rotating pinned buffers on both sides of the GPU: the decoder writes frames straight into a pinned input
            # batch, vex_submit streams it (H2D -> kernels -> D2H on the ctx's three streams) into a pinned output
            # batch, the encoder reads frames straight out of that.  Two batches stay submitted-but-unwaited, so the
            # copies of one overlap the kernels of the other ACROSS calls and the GPU does not drain between batches.
            B, inflight = self.batch_size, 2
            fin = height * width * 3
            slots = pipeline.n_slots(self.pipeline_depth, inflight)
            ins = [self._pinned_buf(f"vin{i}", B * fin).numpy()[: B * fin].reshape(B, height, width, 3) for i in range(slots)]
            outs = [self._pinned_buf(f"vout{i}", B * fin * 4) for i in range(slots)]
            n_read = 0

            def read_batch():
                nonlocal n_read
                view = ins[n_read % slots]
                n = 0
                while n < B:
                    ret, frame = cap.read(view[n])                       # decodes into the pinned row when the size matches
                    if not ret:
                        break
                    if frame.shape != view[n].shape:
                        raise ValueError(f"frame {frame.shape} does not match the container size {(height, width, 3)}")
                    if frame.ctypes.data != view[n].ctypes.data:
                        view[n] = frame
                    n += 1
                if n == 0:
                    return None
                n_read += 1
                return n

            def submit(n, slot):
                t = C.c_ulonglong()
                rc = self._lib.vex_submit(self._ctx, mode, ins[slot].ctypes.data, outs[slot].data_ptr(), n, height, width,
                                          flags, C.byref(t))
                _lib.check(self._ctx, rc, "vex_submit")
                return t.value, slot, n

            def finish(handle):
                t, slot, n = handle
                _lib.check(self._ctx, self._lib.vex_wait(self._ctx, t), "vex_wait")
                return outs[slot].numpy()[: n * fin * 4].reshape(n, 2 * height, 2 * width, 3)

            def write(res):
                nonlocal n_done
                for i in range(res.shape[0]):
                    out.write(res[i])
                n_done += res.shape[0]

            # decode and encode on their own threads, submit / wait on this one
            pipeline.run_ordered_pipeline(read_batch, submit, write, depth=self.pipeline_depth, finish=finish,
                                          inflight=inflight)
        finally:
            cap.release()
            if out is not None:
                out.release()
        return self.video_stats(n_done, width, height, self.scale, self.model_name, float(out_fps),
                                video_io.video_metadata(input_path))       # utils/video_utils.py:38 key set

    @staticmethod
    def video_stats(n_done: int, width: int, height: int, scale: int, model_name: str, out_fps: float, metadata) -> Dict:
        """The statistics dictionary of the file-to-file calls: the keys the agent (video_enhancer_sota.py:202
        `output_frames`), `@track_enhancement_performance` (utils/performance_monitor.py:497-505) and
        `BasicUpscalerFallback.enhance_video` (basic_upscaler_fallback.py:154-162) read or return."""
        return {
            "frames_processed": n_done, "frames_failed": 0, "input_frames": n_done, "output_frames": n_done,
            "scale": scale, "model": f"B200Upscaler-{model_name}",
            "input_resolution": f"{width}x{height}", "output_resolution": f"{width * scale}x{height * scale}",
            "fps": out_fps, "_size": (width, height),
            "input_metadata": metadata,
        }
