"""CPU emulation of the ROUNDING POINTS of the Real-ESRGAN GPU path (csrc/rrdb.cu, csrc/rrdb_tc.cuh) - a test helper, not
the product and not the oracle.  It answers one question for the parity tests: how far from the fp32 oracle does an fp16
network with fp32 accumulation land at all, so that the GPU's own distance can be judged against the number format instead
of against a guessed constant.

Rounding points it reproduces (file:line of the GPU side in brackets):
  * conv_first: fp32 arithmetic (the hi + lo bf16 weight pair carries 16 mantissa bits), output stored fp16  [rrdb.cu conv_first image]
  * every conv3x3 of the trunk and heads: fp16 weights, fp16 activations, fp32 accumulate, fp32 bias, LeakyReLU in fp32,
    ONE rounding to fp16 at the store  [rrdb_tc.cuh epilogue]
  * conv5: ``fmaf(v, 0.2, x)`` and, for rdb3, ``fmaf(., 0.2, x_rrdb)`` in fp32 before that one rounding  [rrdb.cu :385-386]
  * conv_body: ``+ feat`` in fp32, stored fp16; the nearest-2x replicas are copies
  * conv_last: fp16 weights, fp32 accumulate, clamp, x255, truncating cast  [thin_tc.cuh gemm27 MODE_LAST_F16]
The summation ORDER inside a dot product differs from the tensor core's, so the GPU is not expected to equal this bit for
bit: an fp32 sum that lands within 1e-7 of an fp16 tie rounds either way."""
import numpy as np
import torch
import torch.nn.functional as F


def _h(t: torch.Tensor) -> torch.Tensor:
    return t.half().float()                                # round-to-nearest-even to fp16, saturation never reached in the tests


def _conv(x, conv):
    return F.conv2d(x, _h(conv.weight), conv.bias, 1, 1)


def _lrelu(x):
    return F.leaky_relu(x, 0.2)


def _rdb(x, rdb, outer=None):
    x1 = _h(_lrelu(_conv(x, rdb.conv1)))
    x2 = _h(_lrelu(_conv(torch.cat((x, x1), 1), rdb.conv2)))
    x3 = _h(_lrelu(_conv(torch.cat((x, x1, x2), 1), rdb.conv3)))
    x4 = _h(_lrelu(_conv(torch.cat((x, x1, x2, x3), 1), rdb.conv4)))
    v = _conv(torch.cat((x, x1, x2, x3, x4), 1), rdb.conv5) * 0.2 + x
    if outer is not None:
        v = v * 0.2 + outer
    return _h(v)


def forward_f32(model, x: torch.Tensor) -> torch.Tensor:
    """x: fp32 NCHW in [0,1] (bytes / 255).  Returns conv_last's fp32 output before the clamp."""
    with torch.no_grad():
        feat = _h(F.conv2d(x, model.conv_first.weight, model.conv_first.bias, 1, 1))
        h = feat
        for blk in model.body:
            o = _rdb(h, blk.rdb1)
            o = _rdb(o, blk.rdb2)
            h = _rdb(o, blk.rdb3, outer=h)
        feat = _h(_conv(h, model.conv_body) + feat)
        feat = _h(_lrelu(_conv(F.interpolate(feat, scale_factor=2, mode="nearest"), model.conv_up1)))
        if model.scale == 4:
            feat = _h(_lrelu(_conv(F.interpolate(feat, scale_factor=2, mode="nearest"), model.conv_up2)))
        hr = _h(_lrelu(_conv(feat, model.conv_hr)))
        return _conv(hr, model.conv_last)


def process_single(model, img: np.ndarray) -> np.ndarray:
    """Same pre/post steps as oracle/realesrgan_ref.py: process_single (RGB uint8 HWC in and out)."""
    x = torch.from_numpy(img.transpose(2, 0, 1)).float().unsqueeze(0) / 255.0
    out = torch.clamp(forward_f32(model, x), 0, 1).squeeze(0).numpy()
    return (out.transpose(1, 2, 0) * 255).astype(np.uint8)
