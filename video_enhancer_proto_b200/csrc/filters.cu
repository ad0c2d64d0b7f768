// the remaining per-frame stages of BasicUpscalerFallback (models/enhancement/frame/
// basic_upscaler_fallback.py:113-128) as HBM-bound uint8 stencil kernels, so that the whole `frame/` fallback can stay on
// the GPU:  cv2.bilateralFilter(frame, 5, 50, 50)  ->  cv2.resize(..., INTER_LANCZOS4 | INTER_CUBIC)  ->
// cv2.filter2D(upscaled, -1, 0.1 * [[-1,-1,-1],[-1,9,-1],[-1,-1,-1]])  ->  clip, uint8.
// The arithmetic follows OpenCV 4.x (the library the reference delegates to, requirements.txt:18):
//   * Lanczos4, 8-bit: 11-bit fixed-point taps, int32 horizontal then vertical pass, (v + 2^21) >> 22, saturate.  For the
//     exact 2x ratio only two phases exist: t = 0.25 -> {-31,114,-312,1830,579,-188,64,-8}, t = 0.75 -> the mirror
//     (tests/test_filters_gpu.py checks the table against cv2 itself).  Border: replicate.  Bit-exact.
//   * filter2D, 8-bit source, float kernel: fp32 FMA chain over the taps in row-major order, round half to even,
//     saturate; border REFLECT_101.
//   * bilateralFilter, 8-bit 3-channel: radius 2, the 13 taps with i^2 + j^2 <= 4, weight = space[k] *
//     colour[|db| + |dg| + |dr|] (fp32 tables computed in double on the host like OpenCV), fp32 sums, round half to even;
//     border REFLECT_101.
#include <algorithm>
#include <math.h>

#include "vex_internal.cuh"

namespace vex {
namespace {

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
  return i;
}
__device__ __forceinline__ uint8_t sat_u8_rn(float v) {          // cv::saturate_cast<uchar>(cvRound(v))
  return (uint8_t)min(max(__float2int_rn(v), 0), 255);
}

// ------------------------------------------------------------------------------------------------ filter2D 3x3
__global__ void filter3x3_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W,
                                 float k0, float k1, float k2, float k3, float k4, float k5, float k6, float k7, float k8) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= W || y >= H) return;
  const size_t fo = (size_t)blockIdx.z * H * W * 3;
  const float kk[9] = {k0, k1, k2, k3, k4, k5, k6, k7, k8};
  float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int dy = 0; dy < 3; ++dy) {
    const int yy = reflect101(y + dy - 1, H);
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
      const int xx = reflect101(x + dx - 1, W);
      const uint8_t* p = in + fo + ((size_t)yy * W + xx) * 3;
      s0 = fmaf(kk[dy * 3 + dx], (float)p[0], s0);
      s1 = fmaf(kk[dy * 3 + dx], (float)p[1], s1);
      s2 = fmaf(kk[dy * 3 + dx], (float)p[2], s2);
    }
  }
  uint8_t* o = out + fo + ((size_t)y * W + x) * 3;
  o[0] = sat_u8_rn(s0); o[1] = sat_u8_rn(s1); o[2] = sat_u8_rn(s2);
}

// ------------------------------------------------------------------------------------------------ bilateral 5x5
struct BilateralTab {
  float space[13];
  signed char di[13], dj[13];
};

__global__ void bilateral5_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W,
                                  const float* __restrict__ color_w, const BilateralTab tab) {
  __shared__ float cw[768];
  for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < 768; i += blockDim.x * blockDim.y) cw[i] = color_w[i];
  __syncthreads();
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= W || y >= H) return;
  const size_t fo = (size_t)blockIdx.z * H * W * 3;
  const uint8_t* c = in + fo + ((size_t)y * W + x) * 3;
  const int b0 = c[0], g0 = c[1], r0 = c[2];
  float sb = 0.f, sg = 0.f, sr = 0.f, ws = 0.f;
#pragma unroll
  for (int k = 0; k < 13; ++k) {
    const int yy = reflect101(y + tab.di[k], H), xx = reflect101(x + tab.dj[k], W);
    const uint8_t* p = in + fo + ((size_t)yy * W + xx) * 3;
    const int b = p[0], g = p[1], r = p[2];
    const float w = tab.space[k] * cw[abs(b - b0) + abs(g - g0) + abs(r - r0)];
    ws += w;
    sb = fmaf((float)b, w, sb);
    sg = fmaf((float)g, w, sg);
    sr = fmaf((float)r, w, sr);
  }
  const float inv = 1.f / ws;
  uint8_t* o = out + fo + ((size_t)y * W + x) * 3;
  o[0] = sat_u8_rn(sb * inv); o[1] = sat_u8_rn(sg * inv); o[2] = sat_u8_rn(sr * inv);
}

// ------------------------------------------------------------------------------------------------ Lanczos4 2x
constexpr int LZ_TH = 16, LZ_TW = 32;           // input tile; output tile 32 x 64
constexpr int LZ_IN_H = LZ_TH + 8, LZ_IN_W = LZ_TW + 8;
__constant__ int c_lz[2][8] = {{-8, 64, -188, 579, 1830, -312, 114, -31},     // even output (t = 0.75): src j-4 .. j+3
                               {-31, 114, -312, 1830, 579, -188, 64, -8}};    // odd output  (t = 0.25): src j-3 .. j+4

// One block = 16 x 32 input pixels -> 32 x 64 output pixels.  Stage the input tile (+4 halo, replicate border) as bytes,
// run the horizontal pass into int32 shared memory for the 24 rows the vertical pass needs, then the vertical pass.
// int32 arithmetic wraps exactly like the C code it follows.
__global__ void __launch_bounds__(256)
lanczos4_2x_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W) {
  __shared__ uint8_t s_in[LZ_IN_H][LZ_IN_W * 3];
  __shared__ int s_h[LZ_IN_H][2 * LZ_TW * 3];
  const int x0 = blockIdx.x * LZ_TW, y0 = blockIdx.y * LZ_TH;
  const size_t fi = (size_t)blockIdx.z * H * W * 3, fo = (size_t)blockIdx.z * H * W * 12;
  const int tid = threadIdx.x;
  for (int i = tid; i < LZ_IN_H * LZ_IN_W; i += 256) {
    const int r = i / LZ_IN_W, c = i - r * LZ_IN_W;
    const int yy = min(max(y0 + r - 4, 0), H - 1), xx = min(max(x0 + c - 4, 0), W - 1);
    const uint8_t* p = in + fi + ((size_t)yy * W + xx) * 3;
    s_in[r][c * 3 + 0] = p[0]; s_in[r][c * 3 + 1] = p[1]; s_in[r][c * 3 + 2] = p[2];
  }
  __syncthreads();
  // horizontal: out col d = 2j + e of the tile (j in [0,32)); e = 0 uses tile cols j .. j+7, e = 1 uses j+1 .. j+8
  for (int i = tid; i < LZ_IN_H * 2 * LZ_TW * 3; i += 256) {
    const int r = i / (2 * LZ_TW * 3), q = i - r * (2 * LZ_TW * 3);
    const int d = q / 3, ch = q - d * 3;
    const int j = d >> 1, e = d & 1;
    unsigned acc = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc += (unsigned)(c_lz[e][k] * (int)s_in[r][(j + e + k) * 3 + ch]);
    s_h[r][q] = (int)acc;
  }
  __syncthreads();
  // vertical: out row 2i + e of the tile; e = 0 uses tile rows i .. i+7, e = 1 uses i+1 .. i+8.  4 bytes per thread.
  const int row_bytes = 2 * LZ_TW * 3;                       // 192
  const int valid_bytes = min(row_bytes, (2 * W - 2 * x0) * 3);
  const size_t orow = (size_t)2 * W * 3;
  const bool word_ok = ((orow | (size_t)(2 * x0 * 3)) & 3) == 0 && ((reinterpret_cast<uintptr_t>(out) + fo) & 3) == 0;
  for (int i = tid; i < 2 * LZ_TH * (row_bytes / 4); i += 256) {
    const int orow_t = i / (row_bytes / 4), w4 = i - orow_t * (row_bytes / 4);
    const int oy = 2 * y0 + orow_t;
    if (oy >= 2 * H) continue;
    const int ii = orow_t >> 1, e = orow_t & 1;
    uint32_t packed = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int q = w4 * 4 + b;
      unsigned acc = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) acc += (unsigned)(c_lz[e][k] * s_h[ii + e + k][q]);
      const int v = ((int)(acc + (1u << 21))) >> 22;
      packed |= (uint32_t)min(max(v, 0), 255) << (8 * b);
    }
    uint8_t* o = out + fo + (size_t)oy * orow + (size_t)2 * x0 * 3 + w4 * 4;
    if (word_ok && w4 * 4 + 4 <= valid_bytes) {
      *reinterpret_cast<uint32_t*>(o) = packed;
    } else {
      for (int b = 0; b < 4; ++b)
        if (w4 * 4 + b < valid_bytes) o[b] = (uint8_t)(packed >> (8 * b));
    }
  }
}

}  // namespace

int launch_filter3x3(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, const float* k9, cudaStream_t s) {
  if (!in || !out || !k9 || n <= 0 || h <= 0 || w <= 0 || in == out)
    return fail(ctx, VEX_ERR_INVALID, "vex_filter3x3_u8: bad argument");
  dim3 block(32, 8), grid((w + 31) / 32, (h + 7) / 8, n);
  if (grid.y > 65535 || grid.z > 65535) return fail(ctx, VEX_ERR_INVALID, "vex_filter3x3_u8: frame or batch too large");
  filter3x3_kernel<<<grid, block, 0, s>>>(in, out, h, w, k9[0], k9[1], k9[2], k9[3], k9[4], k9[5], k9[6], k9[7], k9[8]);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  return VEX_OK;
}

int launch_bilateral5(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, float sigma_color,
                      float sigma_space, cudaStream_t s) {
  if (!in || !out || n <= 0 || h <= 0 || w <= 0 || in == out || !(sigma_color > 0.f) || !(sigma_space > 0.f))
    return fail(ctx, VEX_ERR_INVALID, "vex_bilateral5_u8: bad argument");
  // tables exactly as cv::bilateralFilter builds them (double exp, cast to float)
  const double gc = -0.5 / ((double)sigma_color * sigma_color), gs = -0.5 / ((double)sigma_space * sigma_space);
  float cw[768];
  for (int i = 0; i < 768; ++i) cw[i] = (float)exp((double)i * i * gc);
  BilateralTab tab;
  int k = 0;
  for (int i = -2; i <= 2; ++i)
    for (int j = -2; j <= 2; ++j) {
      const double r = sqrt((double)i * i + (double)j * j);
      if (r > 2) continue;
      tab.space[k] = (float)exp(r * r * gs);
      tab.di[k] = (signed char)i; tab.dj[k] = (signed char)j;
      ++k;
    }
  if (!ctx->bilateral_cw) VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&ctx->bilateral_cw), sizeof(cw)));
  VEX_CUDA(ctx, cudaMemcpyAsync(ctx->bilateral_cw, cw, sizeof(cw), cudaMemcpyHostToDevice, s));
  VEX_CUDA(ctx, cudaStreamSynchronize(s));                   // cw lives on this stack frame
  dim3 block(32, 8), grid((w + 31) / 32, (h + 7) / 8, n);
  if (grid.y > 65535 || grid.z > 65535) return fail(ctx, VEX_ERR_INVALID, "vex_bilateral5_u8: frame or batch too large");
  bilateral5_kernel<<<grid, block, 0, s>>>(in, out, h, w, ctx->bilateral_cw, tab);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  return VEX_OK;
}

int launch_lanczos4_2x(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, cudaStream_t s) {
  if (!in || !out || n <= 0 || h <= 0 || w <= 0) return fail(ctx, VEX_ERR_INVALID, "vex_lanczos4_2x_u8: bad argument");
  dim3 grid((w + LZ_TW - 1) / LZ_TW, (h + LZ_TH - 1) / LZ_TH, n);
  if (grid.y > 65535 || grid.z > 65535) return fail(ctx, VEX_ERR_INVALID, "vex_lanczos4_2x_u8: frame or batch too large");
  lanczos4_2x_kernel<<<grid, 256, 0, s>>>(in, out, h, w);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  return VEX_OK;
}

}  // namespace vex
