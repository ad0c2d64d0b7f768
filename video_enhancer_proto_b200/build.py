"""Build libvex.so in-tree with nvcc for sm_100a (no torch cpp_extension: the library is a plain
C-ABI shared object loaded with ctypes, so it needs nothing from torch at build or link time)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libvex.so")
SOURCES = ["vex_api.cu", "bicubic.cu", "cnn.cu", "filters.cu", "rrdb.cu"]
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def source_hash() -> str:
    """sha256 over every file the library is compiled from (csrc/*, include/vex.h), the flags and the probe switch.
    build() compiles it into the library as "vex-src:<hash>" (vex_build_info); _lib.load() refuses a library whose
    hash differs from the sources it sits next to."""
    import hashlib
    h = hashlib.sha256()
    files = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")))
    files.append(os.path.join(HERE, "..", "include", "vex.h"))
    for f in files:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    h.update(" ".join(FLAGS).encode())
    h.update(b"probes" if os.environ.get("VEX_DEBUG_PROBES") == "1" else b"")
    return h.hexdigest()


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    data = open(LIB, "rb").read()
    i = data.find(b"vex-src:")
    return i < 0 or data[i + 8:i + 72].decode("ascii", "replace") != source_hash()


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    objs = []
    bdir = os.path.join(HERE, "build")
    os.makedirs(bdir, exist_ok=True)
    procs = []
    defs = [f'-DVEX_SOURCE_HASH="{source_hash()}"']
    if os.environ.get("VEX_DEBUG_PROBES") == "1":          # wrong-result timing probes (tools/thin_probe.py, tools/conv_dbg.py)
        defs.append("-DVEX_DEBUG_PROBES")
    for src in SOURCES:
        obj = os.path.join(bdir, src.replace(".cu", ".o"))
        cmd = [nvcc(), *FLAGS, *(defs if src == "vex_api.cu" else defs[1:]), "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError(f"nvcc failed on {src}")
    cmd = [nvcc(), "-shared", "-o", LIB, *objs, "-lcudart_static", "-ldl", "-lrt", "-lpthread"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
