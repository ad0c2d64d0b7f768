"""CPU check of the layer algebra the CUDA kernels implement (csrc/cnn.cu packing + kernels), against the oracle:

  * eval-mode BN folded into conv weights / bias, the /255 folded into conv1 (app.py:535, 284-291);
  * ConvTranspose2d(3,64,k4,s2,p1) as four 2x2-tap phases == conv3x3 + PixelShuffle(2) (app.py:280, 298);
  * ConvTranspose2d(64,3,k3,s1,p1) as a conv3x3 with the flipped, transposed kernel (app.py:281, 299);
  * the bf16 operand / activation rounding points of the GPU path and the error budget they leave.

In fp32 the restated algebra must agree with the oracle to rounding error; with the bf16 rounding points
emulated it must sit inside the north_star bar (max-abs <= 2/255, PSNR >= 50 dB) for random-init weights.
"""
import numpy as np
import torch
import torch.nn.functional as F

from oracle import upscaler_ref as O
from video_enhancer_proto_b200 import synth

# (ky, dy) per output-row phase and tap index; same table as pack_upconv1 (cnn.cu) / upconv1_kernel
PHASE_TAPS = [[(1, 0), (3, -1)], [(0, 1), (2, 0)]]


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def gpu_algebra(model, frames, bf16):
    sd = model.state_dict()
    q = _bf if bf16 else (lambda x: x)

    def fold(i):
        s = sd[f"bn{i}.weight"] / torch.sqrt(sd[f"bn{i}.running_var"] + 1e-5)
        return s, (sd[f"conv{i}.bias"] - sd[f"bn{i}.running_mean"]) * s + sd[f"bn{i}.bias"]

    def qh(t):                                        # conv1 runs on fp16 operands (thin_tc.cuh MODE_CONV1_H)
        return t.to(torch.float16).to(torch.float32) if bf16 else t

    x = torch.from_numpy(np.stack(frames)).permute(0, 3, 1, 2).float()          # raw u8 values
    s, sh = fold(1)
    x1 = q(F.relu(F.conv2d(x, qh(sd["conv1.weight"] * s[:, None, None, None] / 255.0), sh, padding=1)))
    s, sh = fold(2)
    x2 = q(F.relu(F.conv2d(x1, q(sd["conv2.weight"] * s[:, None, None, None]), sh, padding=1)) + x1)
    s, sh = fold(3)
    x3 = q(F.relu(F.conv2d(x2, q(sd["conv3.weight"] * s[:, None, None, None]), sh, padding=1)))
    s, sh = fold(4)
    x4 = q(F.relu(F.conv2d(x3, q(sd["conv4.weight"] * s[:, None, None, None]), sh, padding=1)))
    x5 = F.conv2d(x4, q(sd["conv5.weight"]), sd["conv5.bias"], padding=1)       # gemm27: bf16 weight image, fp32 x5
    n, _, H, W = x5.shape
    u1w, u1b = q(sd["upconv1.weight"]), sd["upconv1.bias"]                      # tail: bf16 weights, x5 rounded to bf16 as operand
    x5_out = x5
    x5 = q(x5)
    up = torch.zeros(n, u1w.shape[1], 2 * H, 2 * W)
    xp = F.pad(x5, (1, 1, 1, 1))
    for py in range(2):
        for px in range(2):
            acc = u1b[None, :, None, None].expand(n, -1, H, W).clone()
            for ky, dy in PHASE_TAPS[py]:
                for kx, dx in PHASE_TAPS[px]:
                    acc = acc + torch.einsum("nihw,io->nohw", xp[:, :, 1 + dy:1 + dy + H, 1 + dx:1 + dx + W],
                                             u1w[:, :, ky, kx])
            up[:, :, py::2, px::2] = acc
    up = q(F.relu(up))
    out = torch.clamp(F.conv2d(up, q(sd["upconv2.weight"]).flip(2, 3).transpose(0, 1), sd["upconv2.bias"], padding=1), 0, 1)
    return dict(x1=x1, x2=x2, x3=x3, x4=x4, x5=x5_out, up=up, out=out)


def _frames():
    rng = np.random.default_rng(3)
    return [synth.random_frame(40, 56, rng), synth.edges_frame(40, 56, rng)]


def test_fp32_algebra_equals_oracle():
    for model in (O.make_model(0), O.make_stress_model()):
        taps = {}
        with torch.no_grad():
            O.ref_cnn_batch_f32(model, _frames(), taps)
            mine = gpu_algebra(model, _frames(), bf16=False)
        for k, v in mine.items():
            scale = float(taps[k].abs().max())
            assert float((v - taps[k]).abs().max()) <= 2e-5 * max(scale, 1e-3), k


def test_bf16_rounding_points_fit_the_north_star_budget():
    model = O.make_model(0)
    taps = {}
    with torch.no_grad():
        O.ref_cnn_batch_f32(model, _frames(), taps)
        mine = gpu_algebra(model, _frames(), bf16=True)
    err = float((mine["out"] - taps["out"]).abs().max())
    assert err <= 2.0 / 255.0 / 4                      # > 4x headroom under the 2/255 bar
    assert O.psnr(mine["out"].numpy(), taps["out"].numpy()) >= 70.0
