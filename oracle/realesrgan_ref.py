"""CPU restatement of the reference's Real-ESRGAN path.

TEST INFRASTRUCTURE - see ``oracle/__init__.py``.  Follows
``/root/reference/models/enhancement/vsr/realesrgan_handler.py``: ``RRDBBlock`` (:23-36), ``ResidualDenseBlock``
(:38-57), ``RealESRGANNetwork`` (:59-100), and the handler's frame path ``_enhance_frame`` (:286-306),
``_process_single`` (:308-322), ``_tile_process`` (:324-372).  The arithmetic is delegated to the library the reference
delegates to (PyTorch ATen ``conv2d``, ``leaky_relu``, nearest ``interpolate``); this file restates the graph and the host
pre/post steps.  ``oracle/make_golden_rrdb.py`` pins it bit for bit against the reference's own classes, AST-extracted
from that file.  Module and parameter names equal the reference's, so a ``state_dict`` moves either way (the public
``RealESRGAN_x4plus.pth`` uses these names; it cannot be downloaded here, so parity runs on seeded random weights).
"""
from __future__ import annotations

import collections
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


class ResidualDenseBlock(nn.Module):                       # realesrgan_handler.py:38-57
    def __init__(self, num_feat=64, num_grow_ch=32):
        super().__init__()
        self.conv1 = nn.Conv2d(num_feat, num_grow_ch, 3, 1, 1)
        self.conv2 = nn.Conv2d(num_feat + num_grow_ch, num_grow_ch, 3, 1, 1)
        self.conv3 = nn.Conv2d(num_feat + 2 * num_grow_ch, num_grow_ch, 3, 1, 1)
        self.conv4 = nn.Conv2d(num_feat + 3 * num_grow_ch, num_grow_ch, 3, 1, 1)
        self.conv5 = nn.Conv2d(num_feat + 4 * num_grow_ch, num_feat, 3, 1, 1)
        self.lrelu = nn.LeakyReLU(negative_slope=0.2, inplace=True)

    def forward(self, x):
        x1 = self.lrelu(self.conv1(x))
        x2 = self.lrelu(self.conv2(torch.cat((x, x1), 1)))
        x3 = self.lrelu(self.conv3(torch.cat((x, x1, x2), 1)))
        x4 = self.lrelu(self.conv4(torch.cat((x, x1, x2, x3), 1)))
        x5 = self.conv5(torch.cat((x, x1, x2, x3, x4), 1))
        return x5 * 0.2 + x                                # :57


class RRDBBlock(nn.Module):                                # realesrgan_handler.py:23-36
    def __init__(self, num_feat=64, num_grow_ch=32):
        super().__init__()
        self.rdb1 = ResidualDenseBlock(num_feat, num_grow_ch)
        self.rdb2 = ResidualDenseBlock(num_feat, num_grow_ch)
        self.rdb3 = ResidualDenseBlock(num_feat, num_grow_ch)

    def forward(self, x):
        out = self.rdb1(x)
        out = self.rdb2(out)
        out = self.rdb3(out)
        return out * 0.2 + x                               # :36


class RealESRGANNetworkRef(nn.Module):                     # realesrgan_handler.py:59-100
    def __init__(self, num_in_ch=3, num_out_ch=3, scale=4, num_feat=64, num_block=23, num_grow_ch=32):
        super().__init__()
        self.scale = scale
        self.conv_first = nn.Conv2d(num_in_ch, num_feat, 3, 1, 1)
        self.body = nn.Sequential(*[RRDBBlock(num_feat=num_feat, num_grow_ch=num_grow_ch) for _ in range(num_block)])
        self.conv_body = nn.Conv2d(num_feat, num_feat, 3, 1, 1)
        if scale == 4:                                     # :70-74
            self.conv_up1 = nn.Conv2d(num_feat, num_feat, 3, 1, 1)
            self.conv_up2 = nn.Conv2d(num_feat, num_feat, 3, 1, 1)
        elif scale == 2:
            self.conv_up1 = nn.Conv2d(num_feat, num_feat, 3, 1, 1)
        self.conv_hr = nn.Conv2d(num_feat, num_feat, 3, 1, 1)
        self.conv_last = nn.Conv2d(num_feat, num_out_ch, 3, 1, 1)
        self.lrelu = nn.LeakyReLU(negative_slope=0.2, inplace=True)

    def forward(self, x, taps: Optional[dict] = None):
        feat = self.conv_first(x)                          # :87
        if taps is not None:
            taps["first"] = feat
            h = feat
            for i, blk in enumerate(self.body):
                h = blk(h)
                taps[f"body{i}"] = h
            body_feat = self.conv_body(h)
        else:
            body_feat = self.conv_body(self.body(feat))    # :88
        feat = feat + body_feat                            # :89
        if taps is not None:
            taps["trunk"] = feat
        if self.scale == 4:                                # :92-96
            feat = self.lrelu(self.conv_up1(F.interpolate(feat, scale_factor=2, mode="nearest")))
            feat = self.lrelu(self.conv_up2(F.interpolate(feat, scale_factor=2, mode="nearest")))
        elif self.scale == 2:
            feat = self.lrelu(self.conv_up1(F.interpolate(feat, scale_factor=2, mode="nearest")))
        if taps is not None:
            taps["up"] = feat
        hr = self.lrelu(self.conv_hr(feat))
        if taps is not None:
            taps["hr"] = hr
        return self.conv_last(hr)                          # :98-99


def make_network(seed: int = 0, scale: int = 4, num_block: int = 23, gain: float = 1.0) -> RealESRGANNetworkRef:
    """Seeded random-init network (the handler's own fallback when no checkpoint is found, :153-156: torch default
    Kaiming-uniform init).  ``gain`` > 1 scales conv_last so that the output leaves [0,1] and both clamps are hit."""
    torch.manual_seed(seed)
    m = RealESRGANNetworkRef(3, 3, scale, 64, num_block, 32)
    if gain != 1.0:
        with torch.no_grad():
            m.conv_last.weight.mul_(gain)
            m.conv_last.bias.copy_(torch.tensor([0.45, 0.5, 0.55]))
    return m.eval()                                        # :158


def export_state(model: nn.Module) -> "collections.OrderedDict[str, np.ndarray]":
    return collections.OrderedDict((k, v.detach().cpu().float().contiguous().numpy()) for k, v in model.state_dict().items())


def process_single(model: nn.Module, img: np.ndarray) -> np.ndarray:
    """``_process_single`` (:308-322): RGB uint8 HWC -> RGB uint8 (sH, sW, 3); clamp, x255, TRUNCATING cast."""
    img_tensor = torch.from_numpy(img.transpose(2, 0, 1)).float().unsqueeze(0)
    img_tensor = img_tensor / 255.0
    with torch.no_grad():
        output = model(img_tensor)
    output = torch.clamp(output, 0, 1).squeeze(0).cpu().numpy()
    return (output.transpose(1, 2, 0) * 255).astype(np.uint8)


def process_single_f32(model: nn.Module, img: np.ndarray, taps: Optional[dict] = None) -> np.ndarray:
    """Pre-quantisation value of :314-319: clamp(model(x), 0, 1) as fp32 CHW."""
    img_tensor = torch.from_numpy(img.transpose(2, 0, 1)).float().unsqueeze(0) / 255.0
    with torch.no_grad():
        output = model(img_tensor, taps) if taps is not None else model(img_tensor)
    return torch.clamp(output, 0, 1).squeeze(0).numpy()


def tile_plan(h: int, w: int, tile_size: int, tile_pad: int):
    """The tile geometry of ``_tile_process`` (:324-372): for each tile, the padded input window and where the un-padded
    core lands in the output.  Returns [(pad_y0, pad_y1, pad_x0, pad_x1, y0, y1, x0, x1)] in the reference's loop order."""
    tiles_x = (w + tile_size - 1) // tile_size
    tiles_y = (h + tile_size - 1) // tile_size
    plan = []
    for y in range(tiles_y):
        for x in range(tiles_x):
            sx, sy = x * tile_size, y * tile_size
            ex, ey = min(sx + tile_size, w), min(sy + tile_size, h)
            plan.append((max(sy - tile_pad, 0), min(ey + tile_pad, h), max(sx - tile_pad, 0), min(ex + tile_pad, w),
                         sy, ey, sx, ex))
    return plan


def tile_process(model: nn.Module, img: np.ndarray, scale: int, tile_size: int = 512, tile_pad: int = 10) -> np.ndarray:
    h, w = img.shape[:2]
    output = np.zeros((h * scale, w * scale, 3), dtype=np.uint8)
    for (py0, py1, px0, px1, y0, y1, x0, x1) in tile_plan(h, w, tile_size, tile_pad):
        enhanced_tile = process_single(model, img[py0:py1, px0:px1])
        pl, pt = (x0 - px0) * scale, (y0 - py0) * scale
        output[y0 * scale:y1 * scale, x0 * scale:x1 * scale] = \
            enhanced_tile[pt:pt + (y1 - y0) * scale, pl:pl + (x1 - x0) * scale]
    return output


def enhance_frame(model: nn.Module, frame_bgr: np.ndarray, scale: int, tile_size: int = 512, tile_pad: int = 10) -> np.ndarray:
    """``_enhance_frame`` (:286-306): BGR in, BGR out; tiled when either side exceeds tile_size (:292)."""
    frame_rgb = frame_bgr[:, :, ::-1].copy()               # cv2.cvtColor(BGR2RGB) on 3-channel uint8 == channel reversal
    if frame_bgr.shape[0] > tile_size or frame_bgr.shape[1] > tile_size:
        enhanced_rgb = tile_process(model, frame_rgb, scale, tile_size, tile_pad)
    else:
        enhanced_rgb = process_single(model, frame_rgb)
    return enhanced_rgb[:, :, ::-1].copy()


def flops_per_input_pixel(scale: int = 4, num_block: int = 23) -> int:
    rdb = 64 * 32 + 96 * 32 + 128 * 32 + 160 * 32 + 192 * 64
    mac = 3 * 64 + num_block * 3 * rdb + 64 * 64
    if scale == 4:
        mac += 4 * 64 * 64 + 16 * 64 * 64 + 16 * 64 * 64 + 16 * 64 * 3
    else:
        mac += 4 * 64 * 64 + 4 * 64 * 64 + 4 * 64 * 3
    return 2 * 9 * mac
