"""GPU twin of ``RealESRGANHandler`` (models/enhancement/vsr/realesrgan_handler.py:101-387), SURVEY.md §8(f)4.

Same constructor arguments, ``restore_video`` contract and statistics keys (:265-273), the same frame path
(``_enhance_frame`` :286-306, ``_process_single`` :308-322, ``_tile_process`` :324-372) and ``get_model_info`` keys
(:374-387).  The RRDB network (:23-100) runs as hand-written sm_100a kernels behind the C ABI (``vex_rrdb_*``,
include/vex.h): every conv3x3 is a tcgen05 implicit GEMM in fp16 with fp32 accumulation, the dense concatenations are
channel slices of one buffer, the residual scalings and the nearest-neighbour upsampling live in the epilogues.
PyTorch supplies device / pinned buffers and reads checkpoints; there is no CPU fallback (the reference's
``except: return cv2.resize(frame, ...)`` demotion at :303-306 is gone - errors surface).
"""
from __future__ import annotations

import collections
import ctypes as C
import logging
from pathlib import Path
from typing import Dict, Mapping, Optional

import numpy as np
import torch
import torch.nn as nn

from . import _lib, pipeline, video_io

logger = logging.getLogger(__name__)

NUM_FEAT, NUM_GROW_CH = 64, 32            # realesrgan_handler.py:148-150


def state_keys(num_block: int, scale: int):
    """state_dict order of ``RealESRGANNetwork`` (:64-78) == tensor order of ``vex_rrdb_load_weights``."""
    names = ["conv_first"]
    for b in range(num_block):
        for r in (1, 2, 3):
            names += [f"body.{b}.rdb{r}.conv{c}" for c in (1, 2, 3, 4, 5)]
    names += ["conv_body", "conv_up1"] + (["conv_up2"] if scale == 4 else []) + ["conv_hr", "conv_last"]
    return [f"{n}.{p}" for n in names for p in ("weight", "bias")]


def expected_shape(key: str):
    name, p = key.rsplit(".", 1)
    leaf = name.split(".")[-1]
    if leaf == "conv_first":
        co, ci = NUM_FEAT, 3
    elif leaf == "conv_last":
        co, ci = 3, NUM_FEAT
    elif leaf.startswith("conv") and leaf[4:].isdigit() and "rdb" in name:
        c = int(leaf[4:])
        co, ci = (NUM_FEAT if c == 5 else NUM_GROW_CH), NUM_FEAT + (c - 1) * NUM_GROW_CH
    else:
        co, ci = NUM_FEAT, NUM_FEAT
    return (co, ci, 3, 3) if p == "weight" else (co,)


def init_state(seed: Optional[int] = None, scale: int = 4, num_block: int = 23):
    """Random weights by the reference's procedure when no checkpoint is found (:153-156: the module is constructed and
    left at torch's default Conv2d initialisation).  Layers are created in the module's construction order (:64-78), so
    under ``torch.manual_seed(s)`` the tensors are bit-identical to the reference's for the same seed."""
    if seed is not None:
        torch.manual_seed(seed)
    st = collections.OrderedDict()
    convs = collections.OrderedDict()
    convs["conv_first"] = nn.Conv2d(3, NUM_FEAT, 3, 1, 1)
    for b in range(num_block):
        for r in (1, 2, 3):
            for c in (1, 2, 3, 4, 5):
                convs[f"body.{b}.rdb{r}.conv{c}"] = nn.Conv2d(NUM_FEAT + (c - 1) * NUM_GROW_CH,
                                                              NUM_FEAT if c == 5 else NUM_GROW_CH, 3, 1, 1)
    convs["conv_body"] = nn.Conv2d(NUM_FEAT, NUM_FEAT, 3, 1, 1)
    convs["conv_up1"] = nn.Conv2d(NUM_FEAT, NUM_FEAT, 3, 1, 1)
    if scale == 4:
        convs["conv_up2"] = nn.Conv2d(NUM_FEAT, NUM_FEAT, 3, 1, 1)
    convs["conv_hr"] = nn.Conv2d(NUM_FEAT, NUM_FEAT, 3, 1, 1)
    convs["conv_last"] = nn.Conv2d(NUM_FEAT, 3, 3, 1, 1)
    for name, m in convs.items():
        st[f"{name}.weight"] = m.weight.detach().float().contiguous().numpy()
        st[f"{name}.bias"] = m.bias.detach().float().contiguous().numpy()
    return st


def infer_plan(state: Mapping[str, object]):
    blocks = {int(k.split(".")[1]) for k in state if k.startswith("body.")}
    if not blocks:
        raise KeyError("no body.<n>.rdb<k>.conv<j> keys: not a RealESRGANNetwork state_dict")
    return max(blocks) + 1, (4 if "conv_up2.weight" in state else 2)


def normalize_state(state: Mapping[str, object], num_block: int, scale: int):
    out = collections.OrderedDict()
    for k in state_keys(num_block, scale):
        if k not in state:
            raise KeyError(f"state_dict is missing {k}")
        v = state[k]
        a = v.detach().cpu().float().numpy() if isinstance(v, torch.Tensor) else np.asarray(v, dtype=np.float32)
        a = np.ascontiguousarray(a, dtype=np.float32)
        if a.shape != expected_shape(k):
            raise ValueError(f"{k}: shape {a.shape}, expected {expected_shape(k)}")
        out[k] = a
    return out


class B200RealESRGAN:
    """Drop-in for ``RealESRGANHandler``.  NOT thread-safe; one instance per GPU."""

    def __init__(self, model_path: Optional[str] = None, device: str = "cuda", scale: int = 4, tile_size: int = 512,
                 tile_pad: int = 10, state_dict: Optional[Mapping[str, object]] = None, seed: Optional[int] = None,
                 num_block: int = 23):
        if scale not in (2, 4):
            raise ValueError("RealESRGANNetwork has upsampling heads for scale 2 and 4 only (realesrgan_handler.py:69-74)")
        self.device_str = device
        self.scale = scale
        self.tile_size = tile_size
        self.tile_pad = tile_pad
        self.model_path = model_path
        self.num_block = num_block
        self.pipeline_depth = 2
        self._model_loaded = False
        self._pending = (state_dict, seed)
        self._state = None
        self._ctx = C.c_void_p()
        self._lib = None
        self._pinned: Dict[str, torch.Tensor] = {}
        self._dev: Dict[str, torch.Tensor] = {}

    # ------------------------------------------------------------------ lifetime
    def _ordinal(self) -> int:
        d = torch.device(self.device_str if self.device_str != "cuda" else "cuda:0")
        if d.type != "cuda":
            raise RuntimeError(f"B200RealESRGAN needs a CUDA device, got {self.device_str!r} (no CPU fallback)")
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

    def _ensure_model_loaded(self):
        """:134-160: deferred model construction; weights from ``model_path`` when it exists, else random init."""
        if self._model_loaded:
            return
        sd, seed = self._pending
        if sd is None and self.model_path and Path(self.model_path).exists():
            sd = self._read_checkpoint(self.model_path)
        if sd is None:
            logger.warning("No Real-ESRGAN model weights found, using random initialization")   # :156
            sd = init_state(seed, self.scale, self.num_block)
        self.load_state_dict(sd)

    @staticmethod
    def _read_checkpoint(path: str):
        """:190-207: the checkpoint formats the reference accepts."""
        ck = torch.load(path, map_location="cpu", weights_only=True)
        for key in ("params_ema", "params", "state_dict"):
            if key in ck:
                return ck[key]
        return ck

    def load_state_dict(self, state_dict: Mapping[str, object]) -> None:
        self._ensure_ctx()
        num_block, scale = infer_plan(state_dict)
        st = normalize_state(state_dict, num_block, scale)
        desc = _lib.VexRrdbDesc(NUM_FEAT, NUM_GROW_CH, num_block, scale)
        ptrs = (C.c_void_p * len(st))(*[a.ctypes.data for a in st.values()])
        rc = self._lib.vex_rrdb_load_weights(self._ctx, C.byref(desc), ptrs, len(st))
        _lib.check(self._ctx, rc, "vex_rrdb_load_weights")
        self._state, self.num_block, self.scale = st, num_block, scale
        self._model_loaded = True

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
    def _buf(self, pool, tag: str, nbytes: int, pinned: bool) -> torch.Tensor:
        t = pool.get(tag)
        if t is None or t.numel() < nbytes:
            t = (torch.empty(nbytes, dtype=torch.uint8).pin_memory() if pinned
                 else torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{self._ordinal()}"))
            pool[tag] = t
        return t

    def _stream(self):
        return torch.cuda.current_stream(torch.device(f"cuda:{self._ordinal()}")).cuda_stream

    # ------------------------------------------------------------------ device entry points
    def upscale_device(self, frames_u8: torch.Tensor, in_bgr: bool = False, out_bgr: bool = False) -> torch.Tensor:
        """[n,h,w,3] uint8 CUDA tensor -> [n,s*h,s*w,3]: every image in one pass (no tiling), asynchronous."""
        self._ensure_model_loaded()
        x = frames_u8.contiguous()
        n, h, w, _ = x.shape
        out = torch.empty((n, self.scale * h, self.scale * w, 3), dtype=torch.uint8, device=x.device)
        rc = self._lib.vex_rrdb_u8(self._ctx, x.data_ptr(), out.data_ptr(), n, h, w, int(in_bgr), int(out_bgr),
                                   torch.cuda.current_stream(x.device).cuda_stream)
        _lib.check(self._ctx, rc, "vex_rrdb_u8")
        return out

    def frame_device(self, frame_u8: torch.Tensor, in_bgr: bool = False, out_bgr: bool = False) -> torch.Tensor:
        """[h,w,3] uint8 CUDA tensor -> [s*h,s*w,3], tiled like ``_tile_process`` when a side exceeds tile_size."""
        self._ensure_model_loaded()
        x = frame_u8.contiguous()
        h, w, _ = x.shape
        out = torch.empty((self.scale * h, self.scale * w, 3), dtype=torch.uint8, device=x.device)
        rc = self._lib.vex_rrdb_frame_u8(self._ctx, x.data_ptr(), out.data_ptr(), h, w, self.tile_size, self.tile_pad,
                                         int(in_bgr), int(out_bgr), torch.cuda.current_stream(x.device).cuda_stream)
        _lib.check(self._ctx, rc, "vex_rrdb_frame_u8")
        return out

    def debug_f32(self, frames_u8: torch.Tensor, tap: int = 0, in_bgr: bool = False) -> torch.Tensor:
        """Parity hook (vex_rrdb_f32_debug): fp32 NCHW of the selected activation."""
        self._ensure_model_loaded()
        x = frames_u8.contiguous()
        n, h, w, _ = x.shape
        s = self.scale
        if tap == 0:
            shape = (n, 3, s * h, s * w)
        elif tap == 3:
            shape = (n, NUM_FEAT, s * h, s * w)
        elif 1001 <= tap <= 1004:
            shape = (n, NUM_GROW_CH, h, w)
        else:
            shape = (n, NUM_FEAT, h, w)
        out = torch.empty(shape, dtype=torch.float32, device=x.device)
        rc = self._lib.vex_rrdb_f32_debug(self._ctx, x.data_ptr(), out.data_ptr(), n, h, w, int(in_bgr), tap,
                                          torch.cuda.current_stream(x.device).cuda_stream)
        _lib.check(self._ctx, rc, "vex_rrdb_f32_debug")
        return out

    def kernel_launches(self) -> int:
        self._ensure_ctx()
        v = C.c_ulonglong()
        self._lib.vex_kernel_launches(self._ctx, C.byref(v))
        return int(v.value)

    def set_option(self, name: str, value: int) -> int:
        """`vex_set_option` on this handler's context (e.g. ``rrdb_batch``, ``pdl``); returns the previous value."""
        self._ensure_ctx()
        return self._lib.vex_set_option(self._ctx, name.encode(), int(value))

    # ------------------------------------------------------------------ RealESRGANHandler frame path
    def _host_frame(self, frame: np.ndarray, in_bgr: bool, out_bgr: bool, tiled: Optional[bool] = None, slot: Optional[int] = 0) -> np.ndarray:
        """One host frame through the GPU: pinned H2D, the (tiled) network, pinned D2H.  Returns a view of the pinned
        output buffer of `slot` (valid until the next call with that slot); `slot=None` returns a FRESH array the caller
        owns - a pinned block from torch's caching host allocator that the D2H copy lands in directly, instead of a 100 MB
        host memcpy out of a reused buffer (the array keeps the block alive; freed blocks are reused, not re-pinned)."""
        self._ensure_model_loaded()
        if frame.dtype != np.uint8 or frame.ndim != 3 or frame.shape[2] != 3:
            raise ValueError("frames must be uint8 HWC with 3 channels")
        h, w = frame.shape[:2]
        s = self.scale
        fin = h * w * 3
        pin_in = self._buf(self._pinned, "in", fin, True)
        pin_out = (torch.empty(fin * s * s, dtype=torch.uint8, pin_memory=True) if slot is None
                   else self._buf(self._pinned, f"out{slot}", fin * s * s, True))
        d_in = self._buf(self._dev, "in", fin, False)
        d_out = self._buf(self._dev, "out", fin * s * s, False)
        pin_in.numpy()[:fin] = np.ascontiguousarray(frame).ravel()
        d_in[:fin].copy_(pin_in[:fin], non_blocking=True)
        st = self._stream()
        if tiled is None:                                  # the rule of _enhance_frame (:292)
            rc = self._lib.vex_rrdb_frame_u8(self._ctx, d_in.data_ptr(), d_out.data_ptr(), h, w, self.tile_size,
                                             self.tile_pad, int(in_bgr), int(out_bgr), st)
        else:
            rc = self._lib.vex_rrdb_u8(self._ctx, d_in.data_ptr(), d_out.data_ptr(), 1, h, w, int(in_bgr), int(out_bgr), st)
        _lib.check(self._ctx, rc, "vex_rrdb_frame_u8")
        pin_out[: fin * s * s].copy_(d_out[: fin * s * s], non_blocking=True)
        torch.cuda.current_stream(d_in.device).synchronize()
        return pin_out.numpy()[: fin * s * s].reshape(s * h, s * w, 3)

    def _enhance_frame(self, frame: np.ndarray) -> np.ndarray:
        """:286-306: BGR frame in, BGR frame out; tiled when a side exceeds ``tile_size`` (:292).  The two cv2.cvtColor
        calls of the reference are the in_bgr / out_bgr flags of the kernels."""
        return self._host_frame(frame, True, True, None, slot=None)

    def _process_single(self, img: np.ndarray) -> np.ndarray:
        """:308-322: RGB image in, RGB out, the whole image in one pass."""
        return self._host_frame(img, False, False, False, slot=None)

    def _tile_process(self, img: np.ndarray) -> np.ndarray:
        """:324-372: RGB image in, RGB out, tile_size tiles with tile_pad context."""
        self._ensure_model_loaded()
        h, w = img.shape[:2]
        x = torch.from_numpy(np.ascontiguousarray(img)).to(f"cuda:{self._ordinal()}")
        out = torch.empty((self.scale * h, self.scale * w, 3), dtype=torch.uint8, device=x.device)
        # a frame that fits one tile still goes through the tile loop here (one tile, no padding), like the reference's method
        ts = self.tile_size
        if h <= ts and w <= ts:
            rc = self._lib.vex_rrdb_u8(self._ctx, x.data_ptr(), out.data_ptr(), 1, h, w, 0, 0, self._stream())
        else:
            rc = self._lib.vex_rrdb_frame_u8(self._ctx, x.data_ptr(), out.data_ptr(), h, w, ts, self.tile_pad, 0, 0,
                                             self._stream())
        _lib.check(self._ctx, rc, "vex_rrdb_frame_u8")
        return out.cpu().numpy()

    # ------------------------------------------------------------------ handler protocol
    def restore_video(self, input_path: str, output_path: str, **kwargs) -> Dict:
        """:214-284: frame by frame through ``_enhance_frame``; decode and encode on their own threads."""
        import cv2

        self._ensure_model_loaded()
        metadata = video_io.video_metadata(str(input_path))              # utils/video_utils.py:38 (:232-234)
        cap = cv2.VideoCapture(str(input_path))
        if not cap.isOpened():
            raise ValueError(f"Could not open video: {input_path}")
        out = None
        processed = 0
        try:
            fps = metadata.get("fps") or cap.get(cv2.CAP_PROP_FPS)
            width = int(cap.get(cv2.CAP_PROP_FRAME_WIDTH))
            height = int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT))
            total_frames = int(metadata.get("frame_count") or cap.get(cv2.CAP_PROP_FRAME_COUNT))
            out_h, out_w = height * self.scale, width * self.scale
            Path(output_path).parent.mkdir(parents=True, exist_ok=True)
            out = cv2.VideoWriter(str(output_path), cv2.VideoWriter_fourcc(*"mp4v"), fps, (out_w, out_h))   # :238-243
            if not out.isOpened():
                raise ValueError(f"Could not create output video: {output_path}")

            def read():
                ret, frame = cap.read()
                return frame if ret else None

            def process(frame, slot):
                return self._host_frame(frame, True, True, None, slot=slot)

            def write(res):
                nonlocal processed
                out.write(res)
                processed += 1

            pipeline.run_ordered_pipeline(read, process, write, depth=self.pipeline_depth)
        finally:
            cap.release()
            if out is not None:
                out.release()
        return {                                                          # :265-273
            "input_frames": total_frames, "output_frames": processed, "processing_mode": "realesrgan",
            "scale_factor": self.scale, "input_resolution": (height, width), "output_resolution": (out_h, out_w),
            "frames_processed": processed,
        }

    enhance_video = restore_video                                         # the agent may call either name

    def get_model_info(self) -> Dict:
        """:374-387 key set."""
        self._ensure_model_loaded()
        return {
            "name": "Real-ESRGAN",
            "description": "Practical Algorithms for General Image/Video Restoration (B200 tcgen05 kernels)",
            "scale_factor": self.scale,
            "parameters": int(sum(int(np.prod(a.shape)) for a in self._state.values())),
            "device": self.device_str,
            "architecture": "RRDB (Residual in Residual Dense Block)",
            "tile_size": self.tile_size,
            "num_block": self.num_block,
            "precision": "fp16 operands and activations, fp32 accumulate",
        }
