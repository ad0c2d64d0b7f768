"""Weights of the AdvancedUpscaler family for the B200 path.

The reference never loads a checkpoint for this network: ``_create_gpu_upscaler`` builds the module
and runs Xavier-uniform on every conv weight, zero biases, BN at defaults (app.py:303-311).  This
module reproduces that *initialisation procedure* with plain ``torch.nn`` layers constructed in
the reference's order (app.py:273-287), so that under ``torch.manual_seed(s)`` the tensors are
bit-identical to the reference's for the same seed (checked against the oracle in
tests/test_weights.py), and accepts any ``state_dict`` with the reference's key names.
"""
from __future__ import annotations

import collections
from typing import Mapping, Optional, Sequence

import numpy as np
import torch
import torch.nn as nn

# state_dict order == tensor order of vex_cnn_load_weights (include/vex.h)
STATE_DICT_KEYS = (
    "conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "conv3.weight", "conv3.bias",
    "conv4.weight", "conv4.bias", "conv5.weight", "conv5.bias",
    "upconv1.weight", "upconv1.bias", "upconv2.weight", "upconv2.bias",
    "bn1.weight", "bn1.bias", "bn1.running_mean", "bn1.running_var",
    "bn2.weight", "bn2.bias", "bn2.running_mean", "bn2.running_var",
    "bn3.weight", "bn3.bias", "bn3.running_mean", "bn3.running_var",
    "bn4.weight", "bn4.bias", "bn4.running_mean", "bn4.running_var",
)

REFERENCE_WIDTHS = (128, 128, 256, 128)
REFERENCE_UP_CH = 64
EXTRA_PARAMS = ("weight", "bias")
BN_PARAMS = ("weight", "bias", "running_mean", "running_var")


def extra_keys(n_extra: int):
    """Keys of the deepened plans (BASELINE.json configs[4]): block k = conv3x<k> + bn3x<k>, appended after the
    reference's 30 tensors in this order (include/vex.h)."""
    keys = []
    for k in range(1, n_extra + 1):
        keys += [f"conv3x{k}.{p}" for p in EXTRA_PARAMS] + [f"bn3x{k}.{p}" for p in BN_PARAMS]
    return tuple(keys)


def expected_shapes(widths: Sequence[int] = REFERENCE_WIDTHS, up_ch: int = REFERENCE_UP_CH, n_extra: int = 0):
    c1, c2, c3, c4 = widths
    s = collections.OrderedDict()
    for name, (co, ci) in (("conv1", (c1, 3)), ("conv2", (c2, c1)), ("conv3", (c3, c2)),
                           ("conv4", (c4, c3)), ("conv5", (3, c4))):
        s[f"{name}.weight"] = (co, ci, 3, 3)
        s[f"{name}.bias"] = (co,)
    s["upconv1.weight"] = (3, up_ch, 4, 4)      # ConvTranspose2d: (in, out, kh, kw)
    s["upconv1.bias"] = (up_ch,)
    s["upconv2.weight"] = (up_ch, 3, 3, 3)
    s["upconv2.bias"] = (3,)
    for i, c in enumerate((c1, c2, c3, c4), start=1):
        for p in ("weight", "bias", "running_mean", "running_var"):
            s[f"bn{i}.{p}"] = (c,)
    for k in range(1, n_extra + 1):
        s[f"conv3x{k}.weight"] = (c3, c3, 3, 3)
        s[f"conv3x{k}.bias"] = (c3,)
        for p in BN_PARAMS:
            s[f"bn3x{k}.{p}"] = (c3,)
    return collections.OrderedDict((k, s[k]) for k in STATE_DICT_KEYS + extra_keys(n_extra))


def init_state(seed: Optional[int] = None, widths: Sequence[int] = REFERENCE_WIDTHS,
               up_ch: int = REFERENCE_UP_CH, n_extra: int = 0) -> "collections.OrderedDict[str, np.ndarray]":
    """Fresh weights by the reference's procedure (app.py:273-287 construction order, then the
    init loop app.py:305-309).  ``seed=None`` draws from the global RNG like the reference."""
    if seed is not None:
        torch.manual_seed(seed)
    c1, c2, c3, c4 = widths
    layers = collections.OrderedDict()
    layers["conv1"] = nn.Conv2d(3, c1, 3, padding=1)
    layers["conv2"] = nn.Conv2d(c1, c2, 3, padding=1)
    layers["conv3"] = nn.Conv2d(c2, c3, 3, padding=1)
    layers["conv4"] = nn.Conv2d(c3, c4, 3, padding=1)
    layers["conv5"] = nn.Conv2d(c4, 3, 3, padding=1)
    layers["upconv1"] = nn.ConvTranspose2d(3, up_ch, 4, stride=2, padding=1)
    layers["upconv2"] = nn.ConvTranspose2d(up_ch, 3, 3, padding=1)
    for i, c in enumerate((c1, c2, c3, c4), start=1):
        layers[f"bn{i}"] = nn.BatchNorm2d(c)
    for k in range(1, n_extra + 1):                 # deepened plans only; the reference has none
        layers[f"conv3x{k}"] = nn.Conv2d(c3, c3, 3, padding=1)
        layers[f"bn3x{k}"] = nn.BatchNorm2d(c3)
    for m in layers.values():                       # same traversal order as model.modules()
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            nn.init.xavier_uniform_(m.weight)
            nn.init.constant_(m.bias, 0)
    sd = {}
    for name, m in layers.items():
        for k, v in m.state_dict().items():
            sd[f"{name}.{k}"] = v
    return normalize_state(sd, widths, up_ch, n_extra)


def normalize_state(sd: Mapping[str, object], widths: Sequence[int] = REFERENCE_WIDTHS,
                    up_ch: int = REFERENCE_UP_CH, n_extra: int = 0) -> "collections.OrderedDict[str, np.ndarray]":
    """Validate names/shapes and return contiguous fp32 numpy arrays in ABI order."""
    shapes = expected_shapes(widths, up_ch, n_extra)
    out = collections.OrderedDict()
    for k, shp in shapes.items():
        if k not in sd:
            raise KeyError(f"state_dict is missing {k!r}")
        v = sd[k]
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().float().numpy()
        a = np.ascontiguousarray(np.asarray(v, dtype=np.float32))
        if tuple(a.shape) != tuple(shp):
            raise ValueError(f"{k}: expected shape {shp}, got {tuple(a.shape)}")
        out[k] = a
    return out


def infer_plan(sd: Mapping[str, object]):
    """(widths, up_ch, n_extra) from a state_dict's tensor shapes and keys."""
    g = lambda k: tuple(sd[k].shape)
    widths = (g("conv1.weight")[0], g("conv2.weight")[0], g("conv3.weight")[0], g("conv4.weight")[0])
    n_extra = 0
    while f"conv3x{n_extra + 1}.weight" in sd:
        n_extra += 1
    return widths, g("upconv1.weight")[1], n_extra
