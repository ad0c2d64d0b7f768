#!/usr/bin/env python
"""Pin oracle/realesrgan_ref.py to the reference's own Real-ESRGAN code and mint golden vectors.

TEST INFRASTRUCTURE; runs only where /root/reference exists.  ``models/enhancement/vsr/realesrgan_handler.py`` cannot be
imported (huggingface_hub / requests at module level, ``utils`` package side effects), so the three network classes
(:23-100) and the handler's three frame methods (``_enhance_frame``, ``_process_single``, ``_tile_process``, :286-372)
are pulled out of its AST and exec-ed unchanged.  The oracle must equal them bit for bit on seeded weights.

Usage:  python oracle/make_golden_rrdb.py            # regenerate tests/golden/rrdb_cases.npz + manifest entry
        python oracle/make_golden_rrdb.py --check    # only compare oracle vs reference
"""
from __future__ import annotations

import argparse
import ast
import hashlib
import json
import logging
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import realesrgan_ref as R  # noqa: E402
from video_enhancer_proto_b200 import synth  # noqa: E402

REF_PY = "/root/reference/models/enhancement/vsr/realesrgan_handler.py"
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_reference():
    import cv2
    import torch.nn as nn
    import torch.nn.functional as F
    tree = ast.parse(open(REF_PY).read())
    keep = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name in ("RRDBBlock", "ResidualDenseBlock",
                                                                               "RealESRGANNetwork")]
    handler = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "RealESRGANHandler")
    methods = [m for m in handler.body if isinstance(m, ast.FunctionDef) and m.name in ("_enhance_frame", "_process_single",
                                                                                        "_tile_process")]
    shim = ast.ClassDef(name="HandlerFrames", bases=[], keywords=[], body=methods, decorator_list=[])
    mod = ast.Module(body=keep + [shim], type_ignores=[])
    ast.fix_missing_locations(mod)
    g = {"torch": torch, "nn": nn, "F": F, "np": np, "cv2": cv2, "logger": logging.getLogger("reference")}
    exec(compile(mod, REF_PY, "exec"), g)
    return g


def ref_handler(g, model, scale, tile_size, tile_pad):
    h = g["HandlerFrames"]()
    h.model, h.device, h.scale, h.tile_size, h.tile_pad = model, torch.device("cpu"), scale, tile_size, tile_pad
    return h


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    g = load_reference()
    arrays, cases = {}, {}
    # 1. weights and network output: same seed -> same tensors, same forward
    for scale, nb, seed in ((4, 2, 0), (2, 3, 1), (4, 23, 2)):
        torch.manual_seed(seed)
        ref = g["RealESRGANNetwork"](3, 3, scale, 64, nb, 32).eval()
        mine = R.make_network(seed, scale, nb)
        rs, ms = ref.state_dict(), mine.state_dict()
        assert list(rs) == list(ms), "state_dict keys differ"
        for k in rs:
            assert torch.equal(rs[k], ms[k]), k
        x = torch.rand(1, 3, 24, 40, generator=torch.Generator().manual_seed(seed))
        with torch.no_grad():
            assert torch.equal(ref(x), mine(x)), (scale, nb)
        print(f"[ok] RealESRGANNetwork scale={scale} blocks={nb} seed={seed}: weights and forward bit-equal "
              f"({sum(p.numel() for p in ref.parameters())} parameters)")
    # 2. frame path incl. tiling, on small tiles so the tiled branch is exercised cheaply
    for scale, nb, seed, gain, (h, w), tile, pad in ((4, 2, 0, 1.0, (40, 56), 512, 10), (4, 2, 0, 6.0, (40, 56), 512, 10),
                                                     (2, 3, 1, 6.0, (37, 53), 512, 10), (4, 1, 3, 6.0, (70, 45), 32, 10),
                                                     (2, 2, 4, 6.0, (96, 130), 64, 10)):
        mine = R.make_network(seed, scale, nb, gain)
        rng = np.random.default_rng(100 + seed)
        frame = synth.edges_frame(h, w, rng) if seed % 2 else synth.random_frame(h, w, rng)
        hd = ref_handler(g, mine, scale, tile, pad)                       # the reference's methods around the same weights
        want = hd._enhance_frame(frame)
        got = R.enhance_frame(mine, frame, scale, tile, pad)
        assert want.dtype == np.uint8 and np.array_equal(want, got), (scale, nb, h, w, tile)
        key = f"rrdb_s{scale}_b{nb}_seed{seed}_g{gain:g}_{h}x{w}_t{tile}"
        f32 = R.process_single_f32(mine, frame[:, :, ::-1].copy()) if max(h, w) <= tile else None
        arrays[key + "__in"] = frame
        arrays[key + "__out_u8"] = want
        if f32 is not None:
            arrays[key + "__out_f32"] = f32.astype(np.float32)
        cases[key] = {"scale": scale, "num_block": nb, "seed": seed, "gain": gain, "tile_size": tile, "tile_pad": pad,
                      "in_sha256": sha(frame), "out_sha256": sha(want), "out_min": int(want.min()), "out_max": int(want.max())}
        print(f"[ok] _enhance_frame {key}: oracle == reference, output range {want.min()}..{want.max()}")
    if args.check:
        print("check passed")
        return
    np.savez_compressed(os.path.join(GOLDEN, "rrdb_cases.npz"), **arrays)
    with open(os.path.join(GOLDEN, "rrdb_manifest.json"), "w") as f:
        json.dump({"generator": "oracle/make_golden_rrdb.py",
                   "reference": "models/enhancement/vsr/realesrgan_handler.py:23-100, 286-372 (AST-extracted)",
                   "torch": torch.__version__, "cases": cases}, f, indent=1)
    print("wrote", len(cases), "cases")


if __name__ == "__main__":
    main()
