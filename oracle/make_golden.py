#!/usr/bin/env python
"""Mint golden vectors by RUNNING THE REFERENCE'S OWN CODE, and pin the oracle to it.

TEST INFRASTRUCTURE.  Runs only in the build container (needs ``/root/reference``); the GPU
box never executes this — it reads the fixtures committed under ``tests/golden/``.

How the reference is executed: ``/root/reference/app.py`` cannot be imported (gradio,
diffusers, spaces are absent, and import has side effects), so the ``GPUVideoEnhancer``
ClassDef (app.py:242-621) is pulled out of the module's AST and ``exec``-ed unchanged except
for two mechanical edits: the ``@spaces.GPU`` decorators are dropped and the string constant
``'cuda'`` is replaced by ``'cpu'`` (so ``.to('cuda')`` becomes ``.to('cpu')``).  Every other
statement — ``AdvancedUpscaler``, the Xavier init loop, the tensor pre/post-processing, the
bicubic fallback — is the reference's code, byte for byte.

Usage:  python oracle/make_golden.py            # regenerate tests/golden/*
        python oracle/make_golden.py --check    # only compare oracle vs reference
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

from oracle import upscaler_ref as O  # noqa: E402
from video_enhancer_proto_b200 import synth  # noqa: E402

REF_APP = "/root/reference/app.py"
GOLDEN = os.path.join(ROOT, "tests", "golden")


class _CudaToCpu(ast.NodeTransformer):
    def visit_Constant(self, node):
        if node.value == "cuda":
            return ast.copy_location(ast.Constant("cpu"), node)
        return node

    def visit_FunctionDef(self, node):
        node.decorator_list = []
        self.generic_visit(node)
        return node


def load_reference_enhancer():
    """Return an instance of the reference's GPUVideoEnhancer running on CPU."""
    import cv2
    import torch.nn.functional as F

    tree = ast.parse(open(REF_APP).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GPUVideoEnhancer")
    cls = _CudaToCpu().visit(cls)
    mod = ast.Module(body=[cls], type_ignores=[])
    ast.fix_missing_locations(mod)
    g = {
        "torch": torch, "np": np, "cv2": cv2, "F": F, "logger": logging.getLogger("reference"),
        "ZEROGPU_AVAILABLE": True, "spaces": object(), "time": __import__("time"),
        "Path": __import__("pathlib").Path,
    }
    exec(compile(mod, REF_APP, "exec"), g)
    return g["GPUVideoEnhancer"](device="cpu")


def reference_model(enh, seed: int):
    torch.manual_seed(seed)
    enh._gpu_model = enh._create_gpu_upscaler()          # app.py:265-311 verbatim
    enh._gpu_model.eval()                                # app.py:428
    return enh._gpu_model


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def case_inputs(h: int, w: int, seed: int):
    rng = np.random.default_rng(seed)
    return {
        "random": synth.random_frame(h, w, rng),
        "demo": synth.demo_frame(3, size=(w, h), rng=rng, bgr=False) if h >= 64 and w >= 96 else synth.edges_frame(h, w, rng),
        "zeros": synth.constant_frame(h, w, 0),
        "ones": synth.constant_frame(h, w, 255),
        "checker": synth.checker_frame(h, w),
        "impulse": synth.impulse_frame(h, w),
        "edges": synth.edges_frame(h, w, rng),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    os.makedirs(GOLDEN, exist_ok=True)
    enh = load_reference_enhancer()
    manifest = {"generator": "oracle/make_golden.py", "reference": "app.py:242-621 (AST-extracted, 'cuda'->'cpu')",
                "torch": torch.__version__, "cases": {}}

    # ---- 1. weights: reference init == oracle init, bit for bit
    for seed in (0, 1, 2):
        ref = reference_model(enh, seed)
        mine = O.make_model(seed)
        rs, ms = ref.state_dict(), mine.state_dict()
        assert list(rs.keys()) == list(ms.keys()), "state_dict key order differs"
        for k in rs:
            assert torch.equal(rs[k], ms[k]), f"weight mismatch {k} seed {seed}"
        st = O.export_state(mine)
        manifest["cases"][f"weights_seed{seed}"] = {k: sha(v) for k, v in st.items()}
    print("weights: oracle init == reference init (seeds 0,1,2)")

    # ---- 2. CNN forward + pre/post: reference methods vs oracle, bit for bit
    small = {}
    for seed, shapes in ((0, ((37, 53), (64, 96), (120, 160))), (1, ((37, 53),)), (2, ((64, 96),))):
        ref = reference_model(enh, seed)
        mine = O.make_model(seed)
        for (h, w) in shapes:
            for name, frame in case_inputs(h, w, seed * 100 + h).items():
                r_single = enh._enhance_frame_with_gpu(frame)              # app.py:322-352
                r_batch = enh._process_frame_batch_gpu([frame, frame[::-1].copy()])   # app.py:529-553
                o_single = O.ref_cnn_frame(mine, frame)
                o_batch = O.ref_cnn_batch(mine, [frame, frame[::-1].copy()], out_bgr=True)
                assert np.array_equal(r_single, o_single), (seed, h, w, name)
                assert all(np.array_equal(a, b) for a, b in zip(r_batch, o_batch)), (seed, h, w, name)
                key = f"cnn_seed{seed}_{h}x{w}_{name}"
                manifest["cases"][key] = {"in": sha(frame), "out_single_rgb": sha(r_single),
                                          "out_batch_bgr": [sha(x) for x in r_batch]}
                if (h, w) == (37, 53) and name in ("random", "edges", "impulse"):
                    with torch.no_grad():
                        f32 = ref(torch.from_numpy(frame).permute(2, 0, 1).unsqueeze(0).float() / 255.0)[0].numpy()
                    small[f"{key}__in"] = frame
                    small[f"{key}__out_f32"] = f32
                    small[f"{key}__out_u8"] = r_single
    print("cnn: oracle == reference methods on all small cases")

    # stress model (oracle-defined weights; reference GRAPH with those weights loaded)
    stress = O.make_stress_model()
    ref = reference_model(enh, 0)
    ref.load_state_dict(stress.state_dict())
    for (h, w) in ((37, 53), (64, 96)):
        for name, frame in case_inputs(h, w, 777 + h).items():
            r = enh._enhance_frame_with_gpu(frame)
            o = O.ref_cnn_frame(stress, frame)
            assert np.array_equal(r, o), ("stress", h, w, name)
            key = f"cnn_stress_{h}x{w}_{name}"
            manifest["cases"][key] = {"in": sha(frame), "out_single_rgb": sha(r),
                                      "out_min": int(r.min()), "out_max": int(r.max())}
            if (h, w) == (37, 53) and name in ("random", "edges"):
                with torch.no_grad():
                    f32 = ref(torch.from_numpy(frame).permute(2, 0, 1).unsqueeze(0).float() / 255.0)[0].numpy()
                small[f"{key}__in"] = frame
                small[f"{key}__out_f32"] = f32
                small[f"{key}__out_u8"] = r
    manifest["cases"]["weights_stress"] = {k: sha(v) for k, v in O.export_state(stress).items()}
    print("cnn stress weights: oracle == reference graph")

    # ---- 3. bicubic: reference method vs oracle vs numpy restatement vs cv2 twin
    for (h, w) in ((1, 1), (2, 3), (5, 7), (37, 53), (64, 96), (360, 480), (1080, 1920)):
        for name, frame in case_inputs(h, w, 4242 + h).items():
            if (h, w) == (1080, 1920) and name not in ("random", "edges"):
                continue
            r = enh._enhance_frame_cpu_fallback(frame)                      # app.py:359-376
            o = O.ref_bicubic_frame(frame)
            assert np.array_equal(r, o), ("bicubic", h, w, name)
            n = O.np_bicubic2x_u8(frame, 0)
            d = np.abs(n.astype(int) - r.astype(int))
            assert d.max() <= 1, ("np restatement off by >1", h, w, name, int(d.max()))
            entry = {"in": sha(frame), "out_trunc": sha(r), "np_restatement_mismatch_frac": float((d > 0).mean())}
            if h >= 2 and w >= 2:
                c = O.ref_bicubic_cv2(frame)
                n1 = O.np_bicubic2x_u8(frame, 1)
                entry["out_cv2"] = sha(c)
                entry["cv2_vs_np_round_maxdiff"] = int(np.abs(c.astype(int) - n1.astype(int)).max())
            manifest["cases"][f"bicubic_{h}x{w}_{name}"] = entry
            if h <= 37 and name in ("random", "edges", "impulse", "checker"):
                small[f"bicubic_{h}x{w}_{name}__in"] = frame
                small[f"bicubic_{h}x{w}_{name}__out_u8"] = r
    print("bicubic: oracle == reference method; numpy restatement within 1 LSB")

    if args.check:
        return
    np.savez_compressed(os.path.join(GOLDEN, "small_cases.npz"), **small)
    with open(os.path.join(GOLDEN, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    print("wrote", GOLDEN, "(", len(small), "arrays,", len(manifest["cases"]), "cases )")


if __name__ == "__main__":
    main()
