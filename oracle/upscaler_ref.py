"""CPU restatement of the reference's 2x CNN upscaler and bicubic fallback.

TEST INFRASTRUCTURE — see ``oracle/__init__.py``.  The arithmetic is delegated to the
same third-party library the reference delegates to (PyTorch ATen ``conv2d``,
``conv_transpose2d``, ``batch_norm``, ``upsample_bicubic2d``; requirements.txt:9 pins
``torch>=2.4.0,<3.0.0``) plus OpenCV ``resize`` (requirements.txt:18) — this file only
restates the *graph* and the host-side pre/post steps, each citing the reference line it
follows.  ``oracle/make_golden.py`` checks it bit-for-bit against the reference class
AST-extracted from ``/root/reference/app.py``.
"""
from __future__ import annotations

import collections
from typing import Dict, List, Sequence

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

# state_dict order of the reference module (app.py:273-287).  This order is the order of
# the ``tensors`` array of ``vex_cnn_load_weights`` (include/vex.h).
STATE_DICT_KEYS = (
    "conv1.weight", "conv1.bias",
    "conv2.weight", "conv2.bias",
    "conv3.weight", "conv3.bias",
    "conv4.weight", "conv4.bias",
    "conv5.weight", "conv5.bias",
    "upconv1.weight", "upconv1.bias",
    "upconv2.weight", "upconv2.bias",
    "bn1.weight", "bn1.bias", "bn1.running_mean", "bn1.running_var",
    "bn2.weight", "bn2.bias", "bn2.running_mean", "bn2.running_var",
    "bn3.weight", "bn3.bias", "bn3.running_mean", "bn3.running_var",
    "bn4.weight", "bn4.bias", "bn4.running_mean", "bn4.running_var",
)


class AdvancedUpscalerRef(nn.Module):
    """Restatement of ``AdvancedUpscaler`` (app.py:269-301), generalised to a channel plan.

    ``widths=(128, 128, 256, 128)``, ``up_ch=64`` and ``extra=0`` reproduce the reference exactly.  Other
    plans give the "widened/deepened variant" of BASELINE.json configs[4] (same layer kinds): ``extra`` more
    conv3x3 + BatchNorm + ReLU blocks of width c3 after conv3, named conv3x<k> / bn3x<k>.
    """

    def __init__(self, widths: Sequence[int] = (128, 128, 256, 128), up_ch: int = 64, extra: int = 0):
        super().__init__()
        c1, c2, c3, c4 = widths
        self.extra = extra
        assert c1 == c2, "skip connection x2 = relu(bn2(conv2(x1))) + x1 needs equal widths"
        self.conv1 = nn.Conv2d(3, c1, 3, padding=1)        # app.py:273
        self.conv2 = nn.Conv2d(c1, c2, 3, padding=1)       # app.py:274
        self.conv3 = nn.Conv2d(c2, c3, 3, padding=1)       # app.py:275
        self.conv4 = nn.Conv2d(c3, c4, 3, padding=1)       # app.py:276
        self.conv5 = nn.Conv2d(c4, 3, 3, padding=1)        # app.py:277
        self.upconv1 = nn.ConvTranspose2d(3, up_ch, 4, stride=2, padding=1)  # app.py:280
        self.upconv2 = nn.ConvTranspose2d(up_ch, 3, 3, padding=1)            # app.py:281
        self.bn1 = nn.BatchNorm2d(c1)                      # app.py:284
        self.bn2 = nn.BatchNorm2d(c2)                      # app.py:285
        self.bn3 = nn.BatchNorm2d(c3)                      # app.py:286
        self.bn4 = nn.BatchNorm2d(c4)                      # app.py:287
        for k in range(1, extra + 1):                      # not in the reference: configs[4] only
            setattr(self, f"conv3x{k}", nn.Conv2d(c3, c3, 3, padding=1))
            setattr(self, f"bn3x{k}", nn.BatchNorm2d(c3))

    def forward(self, x, taps: Dict[str, torch.Tensor] | None = None):
        x1 = F.relu(self.bn1(self.conv1(x)))               # app.py:291
        x2 = F.relu(self.bn2(self.conv2(x1))) + x1         # app.py:292 (skip AFTER relu)
        x3 = F.relu(self.bn3(self.conv3(x2)))              # app.py:293
        xe, mids = x3, {}
        for k in range(1, self.extra + 1):                 # same block kind as app.py:293
            xe = F.relu(getattr(self, f"bn3x{k}")(getattr(self, f"conv3x{k}")(xe)))
            mids[f"x3e{k}"] = xe
        x4 = F.relu(self.bn4(self.conv4(xe)))              # app.py:294
        x5 = self.conv5(x4)                                # app.py:295
        up = F.relu(self.upconv1(x5))                      # app.py:298
        out = torch.clamp(self.upconv2(up), 0, 1)          # app.py:299
        if taps is not None:
            taps.update(x1=x1, x2=x2, x3=x3, x4=x4, x5=x5, up=up, out=out, **mids)
        return out


def init_reference_weights(model: nn.Module) -> nn.Module:
    """The reference's init loop (app.py:305-309); BN stays at defaults.  Does NOT reseed: the
    reference draws the Xavier values from the RNG stream left by module construction."""
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
    return model


def make_model(seed: int = 0, widths=(128, 128, 256, 128), up_ch: int = 64, extra: int = 0) -> AdvancedUpscalerRef:
    # The reference never seeds (app.py:303-311); parity runs seed once, then construct and
    # run the init loop exactly as ``_create_gpu_upscaler`` does, so that the AST-extracted
    # reference class under the same seed yields bit-identical tensors (oracle/make_golden.py).
    torch.manual_seed(seed)
    m = AdvancedUpscalerRef(widths, up_ch, extra)
    init_reference_weights(m)
    return m.eval()                                        # app.py:331, 428


def make_stress_model(seed: int = 7, gain: float = 1.0) -> AdvancedUpscalerRef:
    """Weights chosen so the output spans [0,1] and hits both clamps (SURVEY.md section 7
    'hard parts': Xavier output range is only ~[0, 0.04]), with non-trivial biases and BN
    statistics so the BN fold (w*g/sqrt(v+eps), (b-mu)*g/sqrt(v+eps)+beta) is exercised."""
    m = make_model(seed)
    g = torch.Generator().manual_seed(seed + 1000)
    with torch.no_grad():
        for name in ("bn1", "bn2", "bn3", "bn4"):
            bn = getattr(m, name)
            c = bn.num_features
            bn.weight.copy_(0.75 + 0.5 * torch.rand(c, generator=g))
            bn.bias.copy_(0.2 * torch.randn(c, generator=g))
            bn.running_mean.copy_(0.1 * torch.randn(c, generator=g))
            bn.running_var.copy_(0.5 + torch.rand(c, generator=g))
        for name in ("conv1", "conv2", "conv3", "conv4", "conv5", "upconv1", "upconv2"):
            conv = getattr(m, name)
            conv.bias.copy_(0.05 * torch.randn(conv.bias.shape, generator=g))
        m.conv5.weight.mul_(4.0 * gain)
        m.upconv1.weight.mul_(3.0)
        m.upconv2.weight.mul_(3.0)
        m.upconv2.bias.copy_(torch.tensor([0.45, 0.5, 0.55]))
    return m.eval()


def export_state(model: nn.Module) -> "collections.OrderedDict[str, np.ndarray]":
    sd = model.state_dict()
    keys = list(STATE_DICT_KEYS)
    for k in range(1, getattr(model, "extra", 0) + 1):
        keys += [f"conv3x{k}.weight", f"conv3x{k}.bias"]
        keys += [f"bn3x{k}.{p}" for p in ("weight", "bias", "running_mean", "running_var")]
    return collections.OrderedDict((k, sd[k].detach().cpu().float().contiguous().numpy()) for k in keys)


# ----------------------------------------------------------------------------- CNN path

def ref_cnn_batch_f32(model: nn.Module, frames_rgb: List[np.ndarray], taps=None) -> torch.Tensor:
    """app.py:533-543 on CPU: uint8 RGB HWC frames -> fp32 NCHW [0,1] output of the model."""
    batch_tensor = []
    for frame in frames_rgb:
        frame_tensor = torch.from_numpy(frame).permute(2, 0, 1).float() / 255.0   # app.py:535
        batch_tensor.append(frame_tensor)
    batch_tensor = torch.stack(batch_tensor)                                       # app.py:539 (no .to('cuda'))
    with torch.no_grad():
        return model(batch_tensor, taps) if taps is not None else model(batch_tensor)  # app.py:542-543


def quantize_u8(x_chw: torch.Tensor) -> np.ndarray:
    """app.py:548-549: permute to HWC, clip(x*255, 0, 255), TRUNCATING uint8 cast."""
    enhanced_frame = x_chw.permute(1, 2, 0).cpu().numpy()
    return np.clip(enhanced_frame * 255, 0, 255).astype(np.uint8)


def ref_cnn_batch(model: nn.Module, frames_rgb: List[np.ndarray], out_bgr: bool = True) -> List[np.ndarray]:
    """``_process_frame_batch_gpu`` (app.py:529-553): RGB uint8 in, BGR uint8 out (2H,2W,3)."""
    enhanced = ref_cnn_batch_f32(model, frames_rgb)
    out = []
    for i in range(enhanced.shape[0]):
        f = quantize_u8(enhanced[i])
        if out_bgr:
            f = f[:, :, ::-1].copy()        # cv2.cvtColor(RGB2BGR) on 3-channel uint8 == channel reversal (app.py:550)
        out.append(f)
    return out


def ref_cnn_frame(model: nn.Module, frame: np.ndarray) -> np.ndarray:
    """``_enhance_frame_with_gpu`` (app.py:334-346): single frame, no colour swap."""
    return ref_cnn_batch(model, [frame], out_bgr=False)[0]


# ------------------------------------------------------------------------- bicubic path

def ref_bicubic_frame(frame: np.ndarray) -> np.ndarray:
    """``_enhance_frame_cpu_fallback`` (app.py:364-376): torch bicubic x2 + clip + truncate."""
    frame_tensor = torch.from_numpy(frame).permute(2, 0, 1).unsqueeze(0).float() / 255.0   # app.py:366
    with torch.no_grad():
        enhanced = F.interpolate(frame_tensor, scale_factor=2, mode="bicubic", align_corners=False)  # app.py:372
    enhanced_frame = enhanced.squeeze(0).permute(1, 2, 0).numpy()                           # app.py:375
    return np.clip(enhanced_frame * 255, 0, 255).astype(np.uint8)                           # app.py:376


def ref_bicubic_frame_f32(frame: np.ndarray) -> np.ndarray:
    """Pre-quantisation value ``enhanced*255`` of app.py:372-376 (float32 HWC)."""
    frame_tensor = torch.from_numpy(frame).permute(2, 0, 1).unsqueeze(0).float() / 255.0
    with torch.no_grad():
        enhanced = F.interpolate(frame_tensor, scale_factor=2, mode="bicubic", align_corners=False)
    return (enhanced.squeeze(0).permute(1, 2, 0).numpy() * 255)


def ref_bicubic_cv2(frame: np.ndarray) -> np.ndarray:
    """OpenCV twin (basic_upscaler_fallback.py:142, realesrgan_fallback.py:104)."""
    import cv2
    h, w = frame.shape[:2]
    return cv2.resize(frame, (w * 2, h * 2), interpolation=cv2.INTER_CUBIC)


# ------------------------------------------------------ numpy restatement of the bicubic

# Keys cubic convolution, A = -0.75 (ATen upsample_bicubic2d / cv2 INTER_CUBIC), evaluated at
# the only two phases a x2 upscale with align_corners=False ever uses:
#   out index d -> src coord s=(d+0.5)/2-0.5, i=floor(s), t=s-i: t=0.75 (d even), 0.25 (d odd)
# taps i-1..i+2, indices clamped to [0, n-1].
W_T025 = np.array([-27.0, 225.0, 67.0, -9.0], dtype=np.float64) / 256.0   # t = 0.25 (d odd)
W_T075 = np.array([-9.0, 67.0, 225.0, -27.0], dtype=np.float64) / 256.0   # t = 0.75 (d even)


def _taps(n: int):
    d = np.arange(2 * n)
    i = np.where(d % 2 == 0, d // 2 - 1, d // 2)             # floor((d+0.5)/2-0.5)
    idx = np.clip(i[:, None] + np.arange(-1, 3)[None, :], 0, n - 1)
    w = np.where((d % 2 == 0)[:, None], W_T075[None, :], W_T025[None, :])
    return idx, w


def np_bicubic2x_f(frame: np.ndarray) -> np.ndarray:
    """Separable restatement (float64) of app.py:366-372 times 255; HWC in, HWC float out."""
    x = frame.astype(np.float64)
    h, w = x.shape[:2]
    iy, wy = _taps(h)
    ix, wx = _taps(w)
    # ATen interpolates along x inside each of the 4 rows, then along y (upsample_bicubic2d).
    rows = np.einsum("hjkc,jk->hjc", x[:, ix, :], wx)          # [h, 2w, c]
    out = np.einsum("ikwc,ik->iwc", rows[iy, :, :], wy)        # [2h, 2w, c]
    return out


def np_bicubic2x_u8(frame: np.ndarray, round_mode: int = 0) -> np.ndarray:
    v = np.clip(np_bicubic2x_f(frame), 0.0, 255.0)
    if round_mode == 0:
        return np.floor(v).astype(np.uint8)                    # truncation (app.py:376)
    return np.floor(v + 0.5).astype(np.uint8)                  # round-to-nearest (cv2)


# -------------------------------------------------------------------------------- metrics

def psnr(a: np.ndarray, b: np.ndarray, peak: float = 1.0) -> float:
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    if mse == 0:
        return float("inf")
    return 10.0 * np.log10(peak * peak / mse)


# ----------------------------------------------------------------------------- BasicUpscalerFallback stages

def ref_basic_upscaler_frame(frame: np.ndarray, model_name: str = "LANCZOS", scale: int = 2, stages: bool = False):
    """The per-frame body of BasicUpscalerFallback.enhance_video (basic_upscaler_fallback.py:113-128), verbatim calls:
    bilateral pre-filter for LANCZOS (:115-117), cv2.resize with the mode's interpolation (:120), the 0.1-scaled
    sharpening filter2D and clip (:123-128).  ``stages=True`` also returns the intermediate images."""
    import cv2
    interpolation = {"LANCZOS": cv2.INTER_LANCZOS4, "CUBIC": cv2.INTER_CUBIC, "LINEAR": cv2.INTER_LINEAR,
                     "AREA": cv2.INTER_AREA, "NEAREST": cv2.INTER_NEAREST}[model_name.upper()]   # :51-57
    out_h, out_w = frame.shape[0] * scale, frame.shape[1] * scale
    den = frame
    if model_name.upper() == "LANCZOS":
        den = cv2.bilateralFilter(frame, 5, 50, 50)
    upscaled0 = cv2.resize(den, (out_w, out_h), interpolation=interpolation)
    kernel = np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]], dtype=np.float32)
    kernel = kernel * 0.1
    upscaled = cv2.filter2D(upscaled0, -1, kernel)
    upscaled = np.clip(upscaled, 0, 255).astype(np.uint8)
    return (upscaled, den, upscaled0) if stages else upscaled


def np_linear2x_u8(frame: np.ndarray) -> np.ndarray:
    """Exact restatement of cv2.resize(frame, (2w, 2h), INTER_LINEAR) for uint8 (basic_upscaler_fallback.py:54,120).
    OpenCV's 8-bit bilinear path is 11-bit fixed point: H pass S = 512*far + 1536*near (phases 0.25 / 0.75, replicate
    border), V pass ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16), then (x + 2) >> 2.  With g = far + 3*near
    (S >> 4 = 32 g) that is ((3 g_near >> 2) + (g_far >> 2) + 2) >> 2: both truncating shifts are OpenCV's."""
    s = frame.astype(np.int64)
    h, w = s.shape[:2]
    dx = np.arange(2 * w)
    k = dx // 2
    far = np.clip(np.where(dx % 2 == 0, k - 1, k + 1), 0, w - 1)
    g = s[:, far] + 3 * s[:, k]
    dy = np.arange(2 * h)
    ky = dy // 2
    fary = np.clip(np.where(dy % 2 == 0, ky - 1, ky + 1), 0, h - 1)
    return (((3 * g[ky]) // 4 + g[fary] // 4 + 2) >> 2).astype(np.uint8)


def np_nearest2x_u8(frame: np.ndarray) -> np.ndarray:
    """cv2.resize INTER_NEAREST (source index floor(d / 2)) and INTER_AREA (enlarging by an integer ratio: OpenCV's
    area-mode bilinear has every fraction 0) at the exact 2x ratio: pixel replication (basic_upscaler_fallback.py:55-56)."""
    return np.repeat(np.repeat(frame, 2, axis=0), 2, axis=1)


LANCZOS4_TAPS_Q11 = {0.25: (-31, 114, -312, 1830, 579, -188, 64, -8), 0.75: (-8, 64, -188, 579, 1830, -312, 114, -31)}


def lanczos4_taps_q11(t: float):
    """OpenCV imgproc ``interpolateLanczos4`` (fp32 coefficients from the sin/cos recurrence table, normalised) followed by
    resize's fixed-point conversion ``saturate_cast<short>(c * 2048)``; the exact 2x ratio only needs t = 0.25, 0.75."""
    x = np.float32(t)
    s45 = 0.70710678118654752440084436210485
    cs = [[1, 0], [-s45, -s45], [0, 1], [s45, -s45], [-1, 0], [s45, s45], [0, -1], [-s45, s45]]
    y0 = -(x + 3) * np.pi * 0.25
    s0, c0 = np.sin(y0), np.cos(y0)
    co = np.zeros(8, np.float32)
    total = np.float32(0)
    for i in range(8):
        y = -(x + 3 - i) * np.pi * 0.25
        co[i] = np.float32((cs[i][0] * s0 + cs[i][1] * c0) / (y * y))
        total = np.float32(total + co[i])
    co = (co * (np.float32(1.0) / total)).astype(np.float32)
    return tuple(int(np.clip(np.rint(np.float32(v) * np.float32(2048)), -32768, 32767)) for v in co)


def np_lanczos4_2x_u8(frame: np.ndarray) -> np.ndarray:
    """Exact restatement of cv2.resize(frame, (2w, 2h), INTER_LANCZOS4) for uint8: horizontal then vertical pass in
    integers with the Q11 taps, replicate border, (v + 2^21) >> 22, saturate.  Small cases only (pure numpy loops)."""
    h, w, _ = frame.shape
    src = frame.astype(np.int64)

    def pass1d(a, n):                       # a: [n, ...] -> [2n, ...] along axis 0
        out = np.zeros((2 * n,) + a.shape[1:], np.int64)
        for d in range(2 * n):
            s = (d + 0.5) * 0.5 - 0.5
            i = int(np.floor(s))
            taps = LANCZOS4_TAPS_Q11[round(s - i, 2)]
            for k in range(8):
                out[d] += taps[k] * a[min(max(i + k - 3, 0), n - 1)]
        return out

    hpass = pass1d(src.transpose(1, 0, 2), w).transpose(1, 0, 2)       # [h, 2w, 3]
    v = pass1d(hpass, h)
    return np.clip((v + (1 << 21)) >> 22, 0, 255).astype(np.uint8)
