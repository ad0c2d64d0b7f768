// The thin layers of AdvancedUpscaler (2.2 % of the FLOPs) as fp32 CUDA-core kernels, plus layout converters
// for the parity taps.  Round-1 versions: correct and coalesced; the tensor-core fusion of the tail
// (conv5 as a 1x1 GEMM in conv4's epilogue, upconv1/upconv2 as PixelShuffle GEMMs) is the next step.
//
//   conv1_u8_kernel      uint8 HWC -> conv1+bn1+relu -> bf16 NHWC      (app.py:291, fuses app.py:535 "/255")
//   conv3x3_to3_kernel   bf16 NHWC (C = 64k) -> 3 channels             (conv5, app.py:295; upconv2+clamp+u8 pack,
//                                                                       app.py:299 + 549)
//   upconv1_kernel       3ch fp32 -> ConvTranspose2d k4 s2 p1 + relu   (app.py:298) == conv3x3 -> PixelShuffle(2)
#pragma once

#include <cuda_bf16.h>

#include "ptx.cuh"

namespace vex {
namespace small {

// ------------------------------------------------------------------------------------------- conv1
// w1g: [27][C1] fp32, k = (ky*3+kx)*3+ci, already multiplied by the BN scale and 1/255.  bias: [C1].
// block (32,8): one output pixel per thread, all C1 channels, 8 at a time.
__global__ void __launch_bounds__(256)
conv1_u8_kernel(const uint8_t* __restrict__ in, __nv_bfloat16* __restrict__ out, const float* __restrict__ w1g,
                const float* __restrict__ bias, int H, int W, int C1, int in_bgr) {
  extern __shared__ float sm_f[];
  float* ws = sm_f;                    // [27][C1]
  float* bs = ws + 27 * C1;            // [C1]
  uint8_t* tile = reinterpret_cast<uint8_t*>(bs + C1);   // [10][34][3]
  const int tid = threadIdx.y * 32 + threadIdx.x;
  const int n = blockIdx.z, x0 = blockIdx.x * 32, y0 = blockIdx.y * 8;
  for (int i = tid; i < 27 * C1; i += 256) ws[i] = w1g[i];
  for (int i = tid; i < C1; i += 256) bs[i] = bias[i];
  const uint8_t* fin = in + (size_t)n * H * W * 3;
  for (int i = tid; i < 10 * 34 * 3; i += 256) {
    const int c = i % 3, px = (i / 3) % 34, py = i / (3 * 34);
    const int x = x0 - 1 + px, y = y0 - 1 + py;
    uint8_t v = 0;                                       // zero padding (Conv2d padding=1)
    if (x >= 0 && x < W && y >= 0 && y < H) v = fin[((size_t)y * W + x) * 3 + (in_bgr ? 2 - c : c)];
    tile[i] = v;
  }
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  float a[27];
#pragma unroll
  for (int ky = 0; ky < 3; ++ky)
#pragma unroll
    for (int kx = 0; kx < 3; ++kx)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        a[(ky * 3 + kx) * 3 + c] = (float)tile[((threadIdx.y + ky) * 34 + threadIdx.x + kx) * 3 + c];
  if (x >= W || y >= H) return;
  __nv_bfloat16* o = out + (((size_t)n * H + y) * W + x) * C1;
  for (int cg = 0; cg < C1; cg += 8) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = bs[cg + j];
#pragma unroll
    for (int k = 0; k < 27; ++k) {
      const float4 wa = *reinterpret_cast<const float4*>(ws + k * C1 + cg);
      const float4 wb = *reinterpret_cast<const float4*>(ws + k * C1 + cg + 4);
      acc[0] = fmaf(a[k], wa.x, acc[0]); acc[1] = fmaf(a[k], wa.y, acc[1]);
      acc[2] = fmaf(a[k], wa.z, acc[2]); acc[3] = fmaf(a[k], wa.w, acc[3]);
      acc[4] = fmaf(a[k], wb.x, acc[4]); acc[5] = fmaf(a[k], wb.y, acc[5]);
      acc[6] = fmaf(a[k], wb.z, acc[6]); acc[7] = fmaf(a[k], wb.w, acc[7]);
    }
    uint4 v;
    v.x = ptx::pack_bf16x2(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f));
    v.y = ptx::pack_bf16x2(fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
    v.z = ptx::pack_bf16x2(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f));
    v.w = ptx::pack_bf16x2(fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f));
    *reinterpret_cast<uint4*>(o + cg) = v;
  }
}
inline size_t conv1_smem(int C1) { return (size_t)(28 * C1) * 4 + 10 * 34 * 3 + 16; }

// ------------------------------------------------------------------------- conv3x3, C (64k) -> 3
// in: bf16 NHWC [n][H][W][C].  wg: [C/64][9][64][4] fp32 (tap = ky*3+kx; 4th output lane is 0).  b: [3].
// MODE 0: out_f4 [n][H][W][4] fp32 = conv + bias                                  (conv5)
// MODE 1: v = clamp(conv + bias, 0, 1); out_u8 [n][H][W][3] = trunc(v*255) (BGR if out_bgr);
//         out_nchw [n][3][H][W] = v when non-null                                 (upconv2 + app.py:299,549)
// block (32,8): tile 32 x 16 pixels, 2 rows per thread; halo tile staged per 64-channel chunk with a
// 144 B pixel pitch (conflict-free 16 B reads across a quarter warp).
constexpr int T3_PIX_PITCH = 144;
constexpr int T3_TILE_H = 16;
constexpr size_t T3_SMEM = (size_t)(T3_TILE_H + 2) * 34 * T3_PIX_PITCH + 9 * 64 * 4 * 4;

template <int MODE>
__global__ void __launch_bounds__(256)
conv3x3_to3_kernel(const __nv_bfloat16* __restrict__ in, const float* __restrict__ wg, const float* __restrict__ b,
                   float* __restrict__ out_f4, uint8_t* __restrict__ out_u8, float* __restrict__ out_nchw,
                   int H, int W, int C, int out_bgr) {
  extern __shared__ __align__(16) uint8_t sm_b[];
  uint8_t* tile = sm_b;                                                    // [18][34][144 B]
  float* ws = reinterpret_cast<float*>(sm_b + (T3_TILE_H + 2) * 34 * T3_PIX_PITCH);   // [9][64][4]
  const int tid = threadIdx.y * 32 + threadIdx.x;
  const int n = blockIdx.z, x0 = blockIdx.x * 32, y0 = blockIdx.y * T3_TILE_H;
  const __nv_bfloat16* fin = in + (size_t)n * H * W * C;
  float acc[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
  for (int c0 = 0; c0 < C; c0 += 64) {
    __syncthreads();
    for (int i = tid; i < (T3_TILE_H + 2) * 34 * 8; i += 256) {
      const int v = i & 7, px = (i >> 3) % 34, py = (i >> 3) / 34;
      const int x = x0 - 1 + px, y = y0 - 1 + py;
      uint4 val = make_uint4(0, 0, 0, 0);
      if (x >= 0 && x < W && y >= 0 && y < H)
        val = __ldg(reinterpret_cast<const uint4*>(fin + ((size_t)y * W + x) * C + c0) + v);
      *reinterpret_cast<uint4*>(tile + (py * 34 + px) * T3_PIX_PITCH + v * 16) = val;
    }
    const float* wsrc = wg + (size_t)(c0 / 64) * 9 * 64 * 4;
    for (int i = tid; i < 9 * 64; i += 256)
      reinterpret_cast<float4*>(ws)[i] = __ldg(reinterpret_cast<const float4*>(wsrc) + i);
    __syncthreads();
#pragma unroll 1
    for (int t = 0; t < 9; ++t) {
      const int ky = t / 3, kx = t % 3;
      const uint8_t* p0 = tile + ((threadIdx.y + ky) * 34 + threadIdx.x + kx) * T3_PIX_PITCH;
      const uint8_t* p1 = p0 + 8 * 34 * T3_PIX_PITCH;
      const float* wt = ws + t * 64 * 4;
#pragma unroll
      for (int v = 0; v < 8; ++v) {
        const uint4 d0 = *reinterpret_cast<const uint4*>(p0 + v * 16);
        const uint4 d1 = *reinterpret_cast<const uint4*>(p1 + v * 16);
        const uint32_t w0[4] = {d0.x, d0.y, d0.z, d0.w}, w1[4] = {d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float4 wl = *reinterpret_cast<const float4*>(wt + (v * 8 + 2 * e) * 4);
          const float4 wh = *reinterpret_cast<const float4*>(wt + (v * 8 + 2 * e + 1) * 4);
          const float a0 = ptx::bf16lo_to_f32(w0[e]), a1 = ptx::bf16hi_to_f32(w0[e]);
          const float c0f = ptx::bf16lo_to_f32(w1[e]), c1f = ptx::bf16hi_to_f32(w1[e]);
          acc[0][0] = fmaf(a0, wl.x, acc[0][0]); acc[0][1] = fmaf(a0, wl.y, acc[0][1]); acc[0][2] = fmaf(a0, wl.z, acc[0][2]);
          acc[0][0] = fmaf(a1, wh.x, acc[0][0]); acc[0][1] = fmaf(a1, wh.y, acc[0][1]); acc[0][2] = fmaf(a1, wh.z, acc[0][2]);
          acc[1][0] = fmaf(c0f, wl.x, acc[1][0]); acc[1][1] = fmaf(c0f, wl.y, acc[1][1]); acc[1][2] = fmaf(c0f, wl.z, acc[1][2]);
          acc[1][0] = fmaf(c1f, wh.x, acc[1][0]); acc[1][1] = fmaf(c1f, wh.y, acc[1][1]); acc[1][2] = fmaf(c1f, wh.z, acc[1][2]);
        }
      }
    }
  }
  const int x = x0 + threadIdx.x;
  if (x >= W) return;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int y = y0 + threadIdx.y + 8 * r;
    if (y >= H) continue;
    const float v0 = acc[r][0] + b[0], v1 = acc[r][1] + b[1], v2 = acc[r][2] + b[2];
    const size_t pix = ((size_t)n * H + y) * W + x;
    if (MODE == 0) {
      reinterpret_cast<float4*>(out_f4)[pix] = make_float4(v0, v1, v2, 0.f);
    } else {
      const float c0f = fminf(fmaxf(v0, 0.f), 1.f), c1f = fminf(fmaxf(v1, 0.f), 1.f), c2f = fminf(fmaxf(v2, 0.f), 1.f);
      if (out_nchw) {
        const size_t plane = (size_t)H * W;
        float* o = out_nchw + (size_t)n * 3 * plane + (size_t)y * W + x;
        o[0] = c0f; o[plane] = c1f; o[2 * plane] = c2f;
      }
      if (out_u8) {
        // np.clip(x*255, 0, 255).astype(np.uint8): fp32 product, truncation (app.py:549)
        const uint8_t q0 = (uint8_t)__float2uint_rz(c0f * 255.0f), q1 = (uint8_t)__float2uint_rz(c1f * 255.0f),
                      q2 = (uint8_t)__float2uint_rz(c2f * 255.0f);
        uint8_t* o = out_u8 + pix * 3;
        o[0] = out_bgr ? q2 : q0; o[1] = q1; o[2] = out_bgr ? q0 : q2;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------ upconv1
// ConvTranspose2d(3, UP, k=4, s=2, p=1) + ReLU.  Output pixel (2y+py, 2x+px) only sees the 2x2 input
// neighbourhood {y+sy(py,a)} x {x+sx(px,b)}, a,b in {0,1}:  py=0: (ky=1, dy=0), (ky=3, dy=-1);
// py=1: (ky=0, dy=+1), (ky=2, dy=0); same in x.
// x5: [n][H][W][4] fp32.  wt: [4 phases][12 = (a*2+b)*3+ci][UP] fp32, bias [UP].  out: bf16 NHWC [n][2H][2W][UP].
// block 128 = 4 warps = 4 phases over 32 consecutive low-res pixels of one row.
__global__ void __launch_bounds__(128)
upconv1_kernel(const float* __restrict__ x5, __nv_bfloat16* __restrict__ out, const float* __restrict__ wt,
               const float* __restrict__ bias, int H, int W, int UP) {
  extern __shared__ float sm_u[];
  float* ws = sm_u;                 // [4][12][UP]
  float* bs = ws + 48 * UP;
  for (int i = threadIdx.x; i < 48 * UP; i += 128) ws[i] = wt[i];
  for (int i = threadIdx.x; i < UP; i += 128) bs[i] = bias[i];
  __syncthreads();
  const int phase = threadIdx.x >> 5, py = phase >> 1, px = phase & 1;
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y, n = blockIdx.z;
  if (x >= W) return;
  float a[12];
#pragma unroll
  for (int ia = 0; ia < 2; ++ia)
#pragma unroll
    for (int ib = 0; ib < 2; ++ib) {
      // (ky, dy) per phase bit and tap index: p=0: {(1,0), (3,-1)}   p=1: {(0,+1), (2,0)}; host packing
      // (pack_upconv1 in cnn.cu) uses the same table.
      const int yy = y + (py ? (ia == 0 ? 1 : 0) : (ia == 0 ? 0 : -1));
      const int xx = x + (px ? (ib == 0 ? 1 : 0) : (ib == 0 ? 0 : -1));
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (yy >= 0 && yy < H && xx >= 0 && xx < W)
        v = __ldg(reinterpret_cast<const float4*>(x5) + ((size_t)n * H + yy) * W + xx);
      a[(ia * 2 + ib) * 3 + 0] = v.x; a[(ia * 2 + ib) * 3 + 1] = v.y; a[(ia * 2 + ib) * 3 + 2] = v.z;
    }
  const float* wp = ws + phase * 12 * UP;
  __nv_bfloat16* o = out + (((size_t)n * 2 * H + 2 * y + py) * (2 * (size_t)W) + 2 * x + px) * UP;
  for (int cg = 0; cg < UP; cg += 8) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = bs[cg + j];
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      const float4 wa = *reinterpret_cast<const float4*>(wp + k * UP + cg);
      const float4 wb = *reinterpret_cast<const float4*>(wp + k * UP + cg + 4);
      acc[0] = fmaf(a[k], wa.x, acc[0]); acc[1] = fmaf(a[k], wa.y, acc[1]);
      acc[2] = fmaf(a[k], wa.z, acc[2]); acc[3] = fmaf(a[k], wa.w, acc[3]);
      acc[4] = fmaf(a[k], wb.x, acc[4]); acc[5] = fmaf(a[k], wb.y, acc[5]);
      acc[6] = fmaf(a[k], wb.z, acc[6]); acc[7] = fmaf(a[k], wb.w, acc[7]);
    }
    uint4 v;
    v.x = ptx::pack_bf16x2(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f));
    v.y = ptx::pack_bf16x2(fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
    v.z = ptx::pack_bf16x2(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f));
    v.w = ptx::pack_bf16x2(fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f));
    *reinterpret_cast<uint4*>(o + cg) = v;
  }
}

// ------------------------------------------------------------------------------- parity-tap converters
// bf16 NHWC [n][H][W][C] -> fp32 NCHW.  One thread per (pixel, 8 channels).
__global__ void nhwc_bf16_to_nchw_f32_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ out,
                                             size_t npix_total, int HW, int C) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int groups = C / 8;
  if (i >= npix_total * groups) return;
  const size_t pix = i / groups;
  const int g = (int)(i % groups);
  const uint4 d = *(reinterpret_cast<const uint4*>(in + pix * C) + g);
  const uint32_t w[4] = {d.x, d.y, d.z, d.w};
  const size_t n = pix / HW, hw = pix % HW;
  float* o = out + (n * C + g * 8) * (size_t)HW + hw;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    o[(size_t)(2 * e) * HW] = ptx::bf16lo_to_f32(w[e]);
    o[(size_t)(2 * e + 1) * HW] = ptx::bf16hi_to_f32(w[e]);
  }
}
// fp32 [n][H][W][4] -> fp32 NCHW [n][3][H][W]
__global__ void nhw4_f32_to_nchw_kernel(const float* __restrict__ in, float* __restrict__ out, size_t npix_total,
                                        int HW) {
  const size_t pix = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix_total) return;
  const float4 v = reinterpret_cast<const float4*>(in)[pix];
  const size_t n = pix / HW, hw = pix % HW;
  float* o = out + n * 3 * (size_t)HW + hw;
  o[0] = v.x; o[HW] = v.y; o[2 * (size_t)HW] = v.z;
}

}  // namespace small
}  // namespace vex
