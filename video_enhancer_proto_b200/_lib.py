"""ctypes binding of libvex.so (include/vex.h).  Fails loudly: no CPU fallback exists."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libvex.so")

VEX_MODE_BICUBIC = 0
VEX_MODE_CNN = 1
VEX_CNN_NUM_TENSORS = 30
VEX_CNN_MAX_EXTRA = 4
VEX_CNN_TENSORS_PER_EXTRA = 6


class VexRrdbDesc(C.Structure):
    _fields_ = [("num_feat", C.c_int), ("num_grow_ch", C.c_int), ("num_block", C.c_int), ("scale", C.c_int)]


class VexCnnDesc(C.Structure):
    _fields_ = [("c1", C.c_int), ("c2", C.c_int), ("c3", C.c_int), ("c4", C.c_int), ("up_ch", C.c_int),
                ("bn_eps", C.c_float), ("n_extra", C.c_int)]


# name -> (restype, argtypes); must list every symbol include/vex.h declares (tests check this)
SIGNATURES = {
    "vex_abi_version": (C.c_int, []),
    "vex_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "vex_destroy": (C.c_int, [C.c_void_p]),
    "vex_last_error": (C.c_char_p, [C.c_void_p]),
    "vex_bicubic2x_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_resize2x_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_lanczos4_2x_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_bilateral5_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p]),
    "vex_filter3x3_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.c_void_p]),
    "vex_cnn_load_weights": (C.c_int, [C.c_void_p, C.POINTER(VexCnnDesc), C.POINTER(C.c_void_p), C.c_int]),
    "vex_cnn2x_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_cnn2x_u8_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_cnn2x_f32_debug": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_cnn_profile_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "vex_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "vex_host_alloc": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "vex_host_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vex_upscale_host": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "vex_submit": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_ulonglong)]),
    "vex_wait": (C.c_int, [C.c_void_p, C.c_ulonglong]),
    "vex_upscale_frames": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "vex_cnn2x_f32in": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_bicubic2x_f32in": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_cnn_profile_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "vex_build_info": (C.c_char_p, []),
    "vex_rrdb_load_weights": (C.c_int, [C.c_void_p, C.POINTER(VexRrdbDesc), C.POINTER(C.c_void_p), C.c_int]),
    "vex_rrdb_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_rrdb_frame_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_rrdb_f32_debug": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "vex_rrdb_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "vex_kernel_launches": (C.c_int, [C.c_void_p, C.POINTER(C.c_ulonglong)]),
    "vex_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
}

_lib = None


def _embedded_hash(path: str):
    """The source hash build.py compiled into the library ("vex-src:<sha256>"), read from the file without loading it."""
    try:
        data = open(path, "rb").read()
    except OSError:
        return None
    i = data.find(b"vex-src:")
    return data[i + 8:i + 72].decode("ascii", "replace") if i >= 0 else None


def load() -> C.CDLL:
    """dlopen libvex.so (built in-tree by ``build.py`` / ``__graft_entry__.build``).  A library whose embedded source
    hash differs from the sources next to it is STALE: it is rebuilt when nvcc is available, otherwise loading fails
    (a prebuilt binary must never silently stand in for different source)."""
    global _lib
    if _lib is None:
        from . import build as _build
        want = _build.source_hash()
        if not os.path.exists(LIB_PATH) or _embedded_hash(LIB_PATH) != want:
            why = "missing" if not os.path.exists(LIB_PATH) else "stale (source hash mismatch)"
            try:
                _build.build(force=True)
            except Exception as e:
                raise RuntimeError(
                    f"{LIB_PATH} is {why} and could not be rebuilt ({e}): run `python -m video_enhancer_proto_b200.build` "
                    "(nvcc, sm_100a).  There is no CPU or PyTorch fallback for this path.") from e
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        info = lib.vex_build_info().decode()
        if want not in info:
            raise RuntimeError(f"{LIB_PATH}: build info {info!r} does not carry the source hash {want}")
        _lib = lib
    return _lib


class VexError(RuntimeError):
    pass


def check(ctx, rc: int, what: str) -> None:
    if rc != 0:
        msg = load().vex_last_error(ctx)
        raise VexError(f"{what} failed ({rc}): {msg.decode() if msg else '?'}")
