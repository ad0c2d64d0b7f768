"""CPU: libvex.so builds for sm_100a, loads, and exports every symbol include/vex.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "vex.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vex_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_header_symbols():
    from video_enhancer_proto_b200 import build, _lib
    path = build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    syms = declared_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in vex.h but not exported"
    assert set(syms) == set(_lib.SIGNATURES), "ctypes signature table out of sync with vex.h"
    assert _lib.load().vex_abi_version() == 2


def test_sass_is_sm100a_native():
    """The CNN kernels must be tcgen05/TMA code, not a recompiled legacy path."""
    import subprocess
    from video_enhancer_proto_b200 import build
    path = build.build()
    sass = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    if "conv3x3_tc_kernel" in sass:
        assert "UTCHMMA" in sass and "UTMALDG" in sass and "LDTM" in sass


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        return
    from video_enhancer_proto_b200 import B200VideoEnhancer, _lib
    import numpy as np
    import pytest
    e = B200VideoEnhancer()
    with pytest.raises(_lib.VexError):
        e.enhance_frame_gpu(np.zeros((4, 4, 3), np.uint8))
    ok, msg = e.process_video("/nonexistent.mp4", "/tmp/x.mp4")
    assert ok is False and msg.startswith("❌")


def test_stale_library_is_refused(tmp_path, monkeypatch):
    """A prebuilt libvex.so compiled from other sources must not load silently: the binding
    compares the hash compiled into the library with the sources next to it."""
    from video_enhancer_proto_b200 import build, _lib
    build.build()
    assert _lib._embedded_hash(_lib.LIB_PATH) == build.source_hash()
    assert build.source_hash() in _lib.load().vex_build_info().decode()
    monkeypatch.setattr(build, "source_hash", lambda: "0" * 64)       # "the sources changed"
    assert build.needs_build()
    monkeypatch.setattr(build, "build", lambda force=False, verbose=False: (_ for _ in ()).throw(RuntimeError("no nvcc")))
    monkeypatch.setattr(_lib, "_lib", None)
    import pytest
    with pytest.raises(RuntimeError, match="stale"):
        _lib.load()


def test_wrong_result_probes_are_gated():
    """vex_set_option refuses the timing probes that skip work unless the library is a VEX_DEBUG_PROBES build; needs a
    context, so it only checks the build-info string without a GPU."""
    from video_enhancer_proto_b200 import _lib
    assert b"probes" not in _lib.load().vex_build_info()
