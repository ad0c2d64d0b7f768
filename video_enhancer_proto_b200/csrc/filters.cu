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

// Fast path (row bytes % 16 == 0, 16-byte aligned buffers, W >= 2): the image is a matrix of bytes with neighbours at
// +-3 bytes; a thread owns one 16-byte column of that matrix and walks down FILT_ROWS rows with a 3-row window of 22
// floats (16 + 3 either side) in registers: per row one 16-byte load, two 4-byte halo loads, 16 x 9 FMAs in exactly the
// row-major tap order of the generic kernel, one 16-byte store.
constexpr int FILT_ROWS = 16;

// uint8 <-> fp32 without the conversion pipe (I2F / F2I issue at a quarter of the FMA rate on B200: 16 per clock per SM,
// tools/ubench/pipes.cu - with 22 + 16 conversions per 16 output bytes they, not the 144 FMAs, bounded this kernel at 0.37
// of HBM).  In: the float 2^23 + b is the bit pattern 0x4B000000 | b (one byte permute), minus 2^23 (exact).  Out:
// clamp to [0, 255], add 1.5 * 2^23 - the sum is rounded to nearest even exactly like cvRound and its low mantissa byte
// IS the integer - and three byte permutes pack four results.  cv::saturate_cast<uchar>(cvRound(v)) clamps after
// rounding; clamping first gives the same byte for every finite v (-0.5 -> 0, 255.5 -> 256 -> 255 either way).
__device__ __forceinline__ float byte_to_f32(uint32_t w, int i) {
  return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540u | i)) - 8388608.0f;
}
__device__ __forceinline__ uint32_t round_sat_bits(float v) {
  return __float_as_uint(fminf(fmaxf(v, 0.f), 255.f) + 12582912.0f);
}
__device__ __forceinline__ uint32_t pack4_low_bytes(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  return __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
}

__device__ __forceinline__ void filt_load_row(const uint8_t* __restrict__ rowp, int col, int row_bytes, float (&f)[22]) {
  const uint4 v = *reinterpret_cast<const uint4*>(rowp + col);
  // halo: bytes col-3..col-1 and col+16..col+18; at the row ends REFLECT_101 in x: pixel -1 -> pixel 1, pixel W -> W-2
  uint32_t lo, hi;
  if (col > 0) lo = *reinterpret_cast<const uint32_t*>(rowp + col - 4) >> 8;            // bytes col-3, col-2, col-1
  else lo = (v.x >> 24) | (v.y << 8);                                                    // bytes 3, 4, 5 (pixel 1)
  if (col + 16 < row_bytes) hi = *reinterpret_cast<const uint32_t*>(rowp + col + 16);    // bytes col+16 .. col+18
  else hi = (v.z >> 16) | (v.w << 16);                                                   // bytes 10, 11, 12 (pixel W-2)
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 3; ++i) f[i] = byte_to_f32(lo, i);
#pragma unroll
  for (int i = 0; i < 16; ++i) f[3 + i] = byte_to_f32(w[i >> 2], i & 3);
#pragma unroll
  for (int i = 0; i < 3; ++i) f[19 + i] = byte_to_f32(hi, i);
}

__global__ void __launch_bounds__(128)
filter3x3_rows_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int row_bytes,
                      float k0, float k1, float k2, float k3, float k4, float k5, float k6, float k7, float k8) {
  const int col = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (col >= row_bytes) return;
  const int y0 = blockIdx.y * FILT_ROWS, y1 = min(y0 + FILT_ROWS, H);
  const size_t fo = (size_t)blockIdx.z * H * row_bytes;
  const uint8_t* base = in + fo;
  const float kk[9] = {k0, k1, k2, k3, k4, k5, k6, k7, k8};
  float r[3][22];                                  // window rows y-1, y, y+1 live in r[t % 3], r[(t+1) % 3], r[(t+2) % 3]
  filt_load_row(base + (size_t)reflect101(y0 - 1, H) * row_bytes, col, row_bytes, r[0]);
  filt_load_row(base + (size_t)y0 * row_bytes, col, row_bytes, r[1]);
#pragma unroll
  for (int t = 0; t < FILT_ROWS; ++t) {            // fully unrolled: the window rotates by renaming, not by moves
    const int y = y0 + t;
    if (y >= y1) break;
    float (&ra)[22] = r[t % 3];
    float (&rb)[22] = r[(t + 1) % 3];
    float (&rc)[22] = r[(t + 2) % 3];
    filt_load_row(base + (size_t)reflect101(y + 1, H) * row_bytes, col, row_bytes, rc);
    uint32_t b[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float a = 0.f;
      a = fmaf(kk[0], ra[i], a); a = fmaf(kk[1], ra[i + 3], a); a = fmaf(kk[2], ra[i + 6], a);
      a = fmaf(kk[3], rb[i], a); a = fmaf(kk[4], rb[i + 3], a); a = fmaf(kk[5], rb[i + 6], a);
      a = fmaf(kk[6], rc[i], a); a = fmaf(kk[7], rc[i + 3], a); a = fmaf(kk[8], rc[i + 6], a);
      b[i] = round_sat_bits(a);
    }
    *reinterpret_cast<uint4*>(out + fo + (size_t)y * row_bytes + col) =
        make_uint4(pack4_low_bytes(b[0], b[1], b[2], b[3]), pack4_low_bytes(b[4], b[5], b[6], b[7]),
                   pack4_low_bytes(b[8], b[9], b[10], b[11]), pack4_low_bytes(b[12], b[13], b[14], b[15]));
  }
}

// Packed variant (row bytes % 32 == 0): one thread owns TWO 16-byte columns, `half` bytes apart, and keeps them in the two
// lanes of fp32x2 registers, so the nine FMAs of an output byte pair are nine FFMA2 instructions.  The kernel above is bound
// by warp-instruction issue (ncu: 68 % issue-active at 18.4 instructions per output byte, FMA pipe 46 %), not by the FMA
// pipe, and FFMA2 has the same FLOP rate with half the instructions: ~10 instructions per byte.  Same fp32 operations in the
// same order per byte, hence the same bytes (tests/test_filters_gpu.py compares both against cv2).  Measured: 0.41 -> 0.43
// of HBM only - halving the FMA instructions did not move it, so issue is not the whole story either; left here.
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_pack(float lo, float hi) {
  f2_t r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(f2_t v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t k, f2_t c) {
  f2_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(k), "l"(c));
  return d;
}
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) {
  f2_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

struct FiltRaw { uint4 v; uint32_t lo, hi; };
__device__ __forceinline__ FiltRaw filt_load_raw(const uint8_t* __restrict__ rowp, int col, int row_bytes) {
  FiltRaw r;
  r.v = *reinterpret_cast<const uint4*>(rowp + col);
  if (col > 0) r.lo = *reinterpret_cast<const uint32_t*>(rowp + col - 4) >> 8;           // bytes col-3 .. col-1
  else r.lo = (r.v.x >> 24) | (r.v.y << 8);                                               // REFLECT_101: pixel 1
  if (col + 16 < row_bytes) r.hi = *reinterpret_cast<const uint32_t*>(rowp + col + 16);   // bytes col+16 .. col+18
  else r.hi = (r.v.z >> 16) | (r.v.w << 16);                                              // pixel W-2
  return r;
}
__device__ __forceinline__ uint32_t filt_byte(const FiltRaw& r, int i) {                 // window byte i of 22 as 0x4B0000bb
  const uint32_t w = i < 3 ? r.lo : i < 7 ? r.v.x : i < 11 ? r.v.y : i < 15 ? r.v.z : i < 19 ? r.v.w : r.hi;
  const int b = i < 3 ? i : i < 19 ? (i - 3) & 3 : i - 19;
  return __byte_perm(w, 0x4B000000u, 0x7540u | b);
}
__device__ __forceinline__ void filt_convert2(const FiltRaw& a, const FiltRaw& b, f2_t (&f)[22]) {
  const f2_t m = f2_pack(-8388608.0f, -8388608.0f);
#pragma unroll
  for (int i = 0; i < 22; ++i)
    f[i] = f2_add(f2_pack(__uint_as_float(filt_byte(a, i)), __uint_as_float(filt_byte(b, i))), m);
}
__device__ __forceinline__ void filt_load_row2(const uint8_t* __restrict__ rowp, int colA, int colB, int row_bytes, f2_t (&f)[22]) {
  const FiltRaw a = filt_load_raw(rowp, colA, row_bytes), b = filt_load_raw(rowp, colB, row_bytes);
  const f2_t m = f2_pack(-8388608.0f, -8388608.0f);
#pragma unroll
  for (int i = 0; i < 22; ++i)
    f[i] = f2_add(f2_pack(__uint_as_float(filt_byte(a, i)), __uint_as_float(filt_byte(b, i))), m);
}

__global__ void __launch_bounds__(128)
filter3x3_rows2_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int row_bytes,
                       float k0, float k1, float k2, float k3, float k4, float k5, float k6, float k7, float k8) {
  const int half = row_bytes / 2;
  const int colA = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (colA >= half) return;
  const int colB = colA + half;
  const int y0 = blockIdx.y * FILT_ROWS, y1 = min(y0 + FILT_ROWS, H);
  const size_t fo = (size_t)blockIdx.z * H * row_bytes;
  const uint8_t* base = in + fo;
  const f2_t kk[9] = {f2_pack(k0, k0), f2_pack(k1, k1), f2_pack(k2, k2), f2_pack(k3, k3), f2_pack(k4, k4),
                      f2_pack(k5, k5), f2_pack(k6, k6), f2_pack(k7, k7), f2_pack(k8, k8)};
  const f2_t zero = f2_pack(0.f, 0.f);
  f2_t r[3][22];
  filt_load_row2(base + (size_t)reflect101(y0 - 1, H) * row_bytes, colA, colB, row_bytes, r[0]);
  filt_load_row2(base + (size_t)y0 * row_bytes, colA, colB, row_bytes, r[1]);
  // (Requesting the raw words of row y + 2 one iteration early was tried: 166 registers, 3 blocks per SM, 0.43 -> 0.30 of HBM.)
#pragma unroll
  for (int t = 0; t < FILT_ROWS; ++t) {
    const int y = y0 + t;
    if (y >= y1) break;
    f2_t (&ra)[22] = r[t % 3];
    f2_t (&rb)[22] = r[(t + 1) % 3];
    f2_t (&rc)[22] = r[(t + 2) % 3];
    filt_load_row2(base + (size_t)reflect101(y + 1, H) * row_bytes, colA, colB, row_bytes, rc);
    uint32_t bA[16], bB[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      f2_t a = zero;
      a = f2_fma(kk[0], ra[i], a); a = f2_fma(kk[1], ra[i + 3], a); a = f2_fma(kk[2], ra[i + 6], a);
      a = f2_fma(kk[3], rb[i], a); a = f2_fma(kk[4], rb[i + 3], a); a = f2_fma(kk[5], rb[i + 6], a);
      a = f2_fma(kk[6], rc[i], a); a = f2_fma(kk[7], rc[i + 3], a); a = f2_fma(kk[8], rc[i + 6], a);
      float lo, hi;
      f2_unpack(a, lo, hi);
      bA[i] = round_sat_bits(lo);
      bB[i] = round_sat_bits(hi);
    }
    uint8_t* orow = out + fo + (size_t)y * row_bytes;
    *reinterpret_cast<uint4*>(orow + colA) =
        make_uint4(pack4_low_bytes(bA[0], bA[1], bA[2], bA[3]), pack4_low_bytes(bA[4], bA[5], bA[6], bA[7]),
                   pack4_low_bytes(bA[8], bA[9], bA[10], bA[11]), pack4_low_bytes(bA[12], bA[13], bA[14], bA[15]));
    *reinterpret_cast<uint4*>(orow + colB) =
        make_uint4(pack4_low_bytes(bB[0], bB[1], bB[2], bB[3]), pack4_low_bytes(bB[4], bB[5], bB[6], bB[7]),
                   pack4_low_bytes(bB[8], bB[9], bB[10], bB[11]), pack4_low_bytes(bB[12], bB[13], bB[14], bB[15]));
  }
}

// ------------------------------------------------------------------------------------------------ bilateral 5x5
struct BilateralTab {
  float space[13];
};
// the 13 taps of a diameter-5 kernel (i^2 + j^2 <= 4) in OpenCV's row-major order
__device__ constexpr int BL_DI[13] = {-2, -1, -1, -1, 0, 0, 0, 0, 0, 1, 1, 1, 2};
__device__ constexpr int BL_DJ[13] = {0, -1, 0, 1, -2, -1, 0, 1, 2, -1, 0, 1, 0};

constexpr int BL_TW = 32, BL_TH = 16;            // pixels per block; the tile is staged with a 2-pixel REFLECT_101 halo

// Pixels are staged as one 32-bit word each (b | g << 8 | r << 16), so a tap is one shared-memory load and the colour
// distance |db| + |dg| + |dr| is one packed sum-of-absolute-differences instruction.
__global__ void __launch_bounds__(BL_TW * BL_TH)
bilateral5_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W,
                  const float* __restrict__ color_w, const BilateralTab tab) {
  __shared__ float cw[768];
  __shared__ uint32_t tile[BL_TH + 4][BL_TW + 4 + 1];
  const int tid = threadIdx.y * BL_TW + threadIdx.x;
  const int x0 = blockIdx.x * BL_TW, y0 = blockIdx.y * BL_TH;
  const size_t fo = (size_t)blockIdx.z * H * W * 3;
  for (int i = tid; i < 768; i += BL_TW * BL_TH) cw[i] = color_w[i];
  for (int i = tid; i < (BL_TH + 4) * (BL_TW + 4); i += BL_TW * BL_TH) {
    const int r = i / (BL_TW + 4), c = i - r * (BL_TW + 4);
    const int yy = reflect101(min(y0 + r - 2, H + 1), H), xx = reflect101(min(x0 + c - 2, W + 1), W);
    const uint8_t* p = in + fo + ((size_t)yy * W + xx) * 3;
    tile[r][c] = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
  }
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  if (x >= W || y >= H) return;
  const uint32_t c0 = tile[threadIdx.y + 2][threadIdx.x + 2];
  float sb = 0.f, sg = 0.f, sr = 0.f, ws = 0.f;
#pragma unroll
  for (int k = 0; k < 13; ++k) {
    const uint32_t p = tile[threadIdx.y + 2 + BL_DI[k]][threadIdx.x + 2 + BL_DJ[k]];
    const float w = tab.space[k] * cw[__vsadu4(p, c0)];
    ws += w;
    // I2F with a byte selector is ONE instruction; this kernel is issue-bound (ncu: issue active 85 %, 382 instructions per
    // pixel), so the two-instruction permute + subtract form that helps filter3x3 costs 9 % here (measured)
    sb = fmaf((float)(p & 0xffu), w, sb);
    sg = fmaf((float)((p >> 8) & 0xffu), w, sg);
    sr = fmaf((float)(p >> 16), w, sr);
  }
  const float inv = 1.f / ws;
  uint8_t* o = out + fo + ((size_t)y * W + x) * 3;
  o[0] = (uint8_t)round_sat_bits(sb * inv); o[1] = (uint8_t)round_sat_bits(sg * inv); o[2] = (uint8_t)round_sat_bits(sr * inv);
}

// ------------------------------------------------------------------------------------------------ Lanczos4 2x
constexpr int LZ_TH = 16, LZ_TW = 32;           // input tile; output tile 32 x 64
constexpr int LZ_IN_H = LZ_TH + 8, LZ_IN_W = LZ_TW + 8;
__constant__ float c_lzf[2][8] = {{-8.f, 64.f, -188.f, 579.f, 1830.f, -312.f, 114.f, -31.f},
                                  {-31.f, 114.f, -312.f, 1830.f, 579.f, -188.f, 64.f, -8.f}};
__constant__ int c_lz[2][8] = {{-8, 64, -188, 579, 1830, -312, 114, -31},     // even output (t = 0.75): src j-4 .. j+3
                               {-31, 114, -312, 1830, 579, -188, 64, -8}};    // odd output  (t = 0.25): src j-3 .. j+4

// One block = 16 x 32 input pixels -> 32 x 64 output pixels.  Stage the input tile (+4 halo, replicate border) as bytes,
// run the horizontal pass into int32 shared memory for the 24 rows the vertical pass needs (in fp32 FMAs: every partial sum
// is an integer below 2^24, so it is exact and twice the integer-multiply rate), then the vertical pass in int32, which
// wraps exactly like the C code it follows.
__global__ void __launch_bounds__(256)
lanczos4_2x_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W) {
  __shared__ __align__(16) uint8_t s_in[LZ_IN_H][LZ_IN_W * 3];
  __shared__ __align__(16) int s_h[LZ_IN_H][2 * LZ_TW * 3];
  const int x0 = blockIdx.x * LZ_TW, y0 = blockIdx.y * LZ_TH;
  const size_t fi = (size_t)blockIdx.z * H * W * 3, fo = (size_t)blockIdx.z * H * W * 12;
  const int tid = threadIdx.x;
  // interior tiles (no clamping in x, 4-byte aligned rows): the 120-byte row segment arrives as 30 words per row
  const bool interior_x = x0 >= 4 && x0 + LZ_TW + 4 <= W && (W & 3) == 0 && ((reinterpret_cast<uintptr_t>(in) + fi) & 3) == 0;
  if (interior_x) {
    for (int i = tid; i < LZ_IN_H * (LZ_IN_W * 3 / 4); i += 256) {
      const int r = i / (LZ_IN_W * 3 / 4), c4 = i - r * (LZ_IN_W * 3 / 4);
      const int yy = min(max(y0 + r - 4, 0), H - 1);
      const uint32_t v = *reinterpret_cast<const uint32_t*>(in + fi + ((size_t)yy * W + (x0 - 4)) * 3 + 4 * c4);
      *reinterpret_cast<uint32_t*>(&s_in[r][4 * c4]) = v;
    }
  } else {
    for (int i = tid; i < LZ_IN_H * LZ_IN_W; i += 256) {
      const int r = i / LZ_IN_W, c = i - r * LZ_IN_W;
      const int yy = min(max(y0 + r - 4, 0), H - 1), xx = min(max(x0 + c - 4, 0), W - 1);
      const uint8_t* p = in + fi + ((size_t)yy * W + xx) * 3;
      s_in[r][c * 3 + 0] = p[0]; s_in[r][c * 3 + 1] = p[1]; s_in[r][c * 3 + 2] = p[2];
    }
  }
  __syncthreads();
  // horizontal: out col d = 2j + e of the tile (j in [0,32)); e = 0 uses tile cols j .. j+7, e = 1 uses j+1 .. j+8.
  // One item = (row, channel, 4 input px): 12 bytes in registers feed 8 outputs.
  for (int i = tid; i < LZ_IN_H * 3 * (LZ_TW / 4); i += 256) {
    const int ch = i % 3, g = (i / 3) % (LZ_TW / 4), r = i / (3 * (LZ_TW / 4));
    float v[12];
#pragma unroll
    for (int c = 0; c < 12; ++c)                              // byte -> float off the conversion pipe: 2^23 + b, minus 2^23
      v[c] = __uint_as_float(0x4B000000u | (uint32_t)s_in[r][(4 * g + c) * 3 + ch]) - 8388608.0f;
#pragma unroll
    for (int d = 0; d < 8; ++d) {
      const int j = d >> 1, e = d & 1;
      float acc = 0.f;                                         // |partial sums| <= 255 * 3126 < 2^24: exact in fp32
#pragma unroll
      for (int k = 0; k < 8; ++k) acc = fmaf(c_lzf[e][k], v[j + e + k], acc);
      // acc is an integer of magnitude < 2^22: acc + 1.5 * 2^23 holds it in the low mantissa bits (no F2I)
      s_h[r][(8 * g + d) * 3 + ch] = (int)(__float_as_uint(acc + 12582912.0f) - 0x4B400000u);
    }
  }
  __syncthreads();
  // vertical: out row 2i + e of the tile; e = 0 uses tile rows i .. i+7, e = 1 uses i+1 .. i+8.  One thread = 4 adjacent
  // output bytes x 8 output rows: 12 rows of 4 ints in registers feed 32 outputs.
  constexpr int row_bytes = 2 * LZ_TW * 3;                   // 192
  const int valid_bytes = min(row_bytes, (2 * W - 2 * x0) * 3);
  const size_t orow = (size_t)2 * W * 3;
  const bool word_ok = ((orow | (size_t)(2 * x0 * 3)) & 3) == 0 && ((reinterpret_cast<uintptr_t>(out) + fo) & 3) == 0;
  if (tid < 4 * (row_bytes / 4)) {
    const int w4 = tid % (row_bytes / 4), rq = tid / (row_bytes / 4);
    int4 hv[12];
#pragma unroll
    for (int r = 0; r < 12; ++r) hv[r] = *reinterpret_cast<const int4*>(&s_h[4 * rq + r][4 * w4]);
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int orow_t = 8 * rq + t, oy = 2 * y0 + orow_t;
      const int ii = t >> 1, e = t & 1;
      unsigned a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int4 h4 = hv[ii + e + k];
        a0 += (unsigned)(c_lz[e][k] * h4.x); a1 += (unsigned)(c_lz[e][k] * h4.y);
        a2 += (unsigned)(c_lz[e][k] * h4.z); a3 += (unsigned)(c_lz[e][k] * h4.w);
      }
      const int v0 = ((int)(a0 + (1u << 21))) >> 22, v1 = ((int)(a1 + (1u << 21))) >> 22,
                v2 = ((int)(a2 + (1u << 21))) >> 22, v3 = ((int)(a3 + (1u << 21))) >> 22;
      const uint32_t packed = (uint32_t)min(max(v0, 0), 255) | ((uint32_t)min(max(v1, 0), 255) << 8) |
                              ((uint32_t)min(max(v2, 0), 255) << 16) | ((uint32_t)min(max(v3, 0), 255) << 24);
      if (oy < 2 * H) {
        uint8_t* o = out + fo + (size_t)oy * orow + (size_t)2 * x0 * 3 + w4 * 4;
        if (word_ok && w4 * 4 + 4 <= valid_bytes) {
          *reinterpret_cast<uint32_t*>(o) = packed;
        } else {
          for (int bb = 0; bb < 4; ++bb)
            if (w4 * 4 + bb < valid_bytes) o[bb] = (uint8_t)(packed >> (8 * bb));
        }
      }
    }
  }
}

}  // namespace

int launch_filter3x3(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, const float* k9, cudaStream_t s) {
  if (!in || !out || !k9 || n <= 0 || h <= 0 || w <= 0 || in == out)
    return fail(ctx, VEX_ERR_INVALID, "vex_filter3x3_u8: bad argument");
  const int row_bytes = w * 3;
  const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  if (row_bytes % 32 == 0 && aligned && w >= 16 && !ctx->opt_force_generic_bicubic && !ctx->opt_force_generic_resize) {
    dim3 grid((row_bytes / 32 + 127) / 128, (h + FILT_ROWS - 1) / FILT_ROWS, n);   // two 16-byte columns per thread, fp32x2
    if (grid.y > 65535 || grid.z > 65535) return fail(ctx, VEX_ERR_INVALID, "vex_filter3x3_u8: frame or batch too large");
    filter3x3_rows2_kernel<<<grid, 128, 0, s>>>(in, out, h, row_bytes, k9[0], k9[1], k9[2], k9[3], k9[4], k9[5], k9[6], k9[7],
                                                k9[8]);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    return VEX_OK;
  }
  if (row_bytes % 16 == 0 && aligned && w >= 8 && !ctx->opt_force_generic_bicubic) {
    dim3 grid((row_bytes / 16 + 127) / 128, (h + FILT_ROWS - 1) / FILT_ROWS, n);
    if (grid.y > 65535 || grid.z > 65535) return fail(ctx, VEX_ERR_INVALID, "vex_filter3x3_u8: frame or batch too large");
    filter3x3_rows_kernel<<<grid, 128, 0, s>>>(in, out, h, row_bytes, k9[0], k9[1], k9[2], k9[3], k9[4], k9[5], k9[6], k9[7],
                                               k9[8]);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    return VEX_OK;
  }
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
  BilateralTab tab;
  int k = 0;
  for (int i = -2; i <= 2; ++i)
    for (int j = -2; j <= 2; ++j) {
      const double r = sqrt((double)i * i + (double)j * j);
      if (r > 2) continue;
      tab.space[k++] = (float)exp(r * r * gs);
    }
  if (!ctx->bilateral_cw || ctx->bilateral_sigma != sigma_color) {     // colour table: uploaded once per sigma
    float cw[768];
    for (int i = 0; i < 768; ++i) cw[i] = (float)exp((double)i * i * gc);
    if (!ctx->bilateral_cw) VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&ctx->bilateral_cw), sizeof(cw)));
    VEX_CUDA(ctx, cudaDeviceSynchronize());                            // no kernel may still be reading the old table
    VEX_CUDA(ctx, cudaMemcpy(ctx->bilateral_cw, cw, sizeof(cw), cudaMemcpyHostToDevice));
    ctx->bilateral_sigma = sigma_color;
  }
  dim3 block(BL_TW, BL_TH), grid((w + BL_TW - 1) / BL_TW, (h + BL_TH - 1) / BL_TH, n);
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

// ---------------------------------------------------------------------------------------------------------------
// cv2.resize(frame, (2w, 2h), INTER_LINEAR / INTER_NEAREST / INTER_AREA) for 8-bit images
// (basic_upscaler_fallback.py:52-57,120).  At the exact ratio 2 OpenCV's 11-bit fixed-point bilinear path reduces to
//   g(r, dx)  = src[r][far(dx)] + 3 * src[r][near(dx)]            (H pass: (512*far + 1536*near) >> 4 = 32 g)
//   out(dy,dx) = ((3 * g(near(dy)) >> 2) + (g(far(dy)) >> 2) + 2) >> 2   (V pass: mul-high by 1536 / 512, rounding shift)
// with near(d) = d >> 1, far(d) = near -+ 1 (d even / odd) clamped to the image (replicate border); the two truncating
// shifts are OpenCV's.  INTER_AREA enlarging by an integer ratio has all fractions 0 and INTER_NEAREST takes floor(d / 2):
// both replicate pixels.  One thread = one 4-byte output word column of RZ_ROWS source rows (a sliding window of g).
// ---------------------------------------------------------------------------------------------------------------
constexpr int RZ_ROWS = 8;
constexpr int RZ_THREADS = 128;

template <bool LINEAR, bool WORDS>
__global__ void __launch_bounds__(RZ_THREADS) resize2x_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                              int h, int w) {
  const int row_in = 3 * w, row_out = 6 * w;
  const int j = blockIdx.x * RZ_THREADS + threadIdx.x;            // output word column
  if (4 * j >= row_out) return;
  const int y0 = blockIdx.y * RZ_ROWS, y1 = min(h, y0 + RZ_ROWS);
  const uint8_t* src = in + (size_t)blockIdx.z * h * row_in;
  uint8_t* dst = out + (size_t)blockIdx.z * 2 * h * row_out;

  int near_b[4], far_b[4], nvalid = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ob = min(4 * j + i, row_out - 1);
    nvalid += (4 * j + i < row_out);
    const int dx = ob / 3, c = ob - 3 * dx, k = dx >> 1;
    const int kf = (dx & 1) ? min(k + 1, w - 1) : max(k - 1, 0);
    near_b[i] = 3 * k + c;
    far_b[i] = 3 * kf + c;
  }
  auto emit = [&](int dy, const int (&v)[4]) {
    uint8_t* d = dst + (size_t)dy * row_out + 4 * j;
    if (WORDS) {
      *reinterpret_cast<uint32_t*>(d) = (uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16) | ((uint32_t)v[3] << 24);
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (i < nvalid) d[i] = (uint8_t)v[i];
    }
  };
  if (!LINEAR) {
    for (int y = y0; y < y1; ++y) {
      const uint8_t* r = src + (size_t)y * row_in;
      int v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = __ldg(r + near_b[i]);
      emit(2 * y, v);
      emit(2 * y + 1, v);
    }
    return;
  }
  auto grow = [&](int y, int (&g)[4]) {
    const uint8_t* r = src + (size_t)y * row_in;
#pragma unroll
    for (int i = 0; i < 4; ++i) g[i] = (int)__ldg(r + far_b[i]) + 3 * (int)__ldg(r + near_b[i]);
  };
  int gp[4], gc[4], gn[4];
  grow(max(y0 - 1, 0), gp);
  grow(y0, gc);
  for (int y = y0; y < y1; ++y) {
    grow(min(y + 1, h - 1), gn);
    int top[4], bot[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = (3 * gc[i]) >> 2;
      top[i] = (m + (gp[i] >> 2) + 2) >> 2;
      bot[i] = (m + (gn[i] >> 2) + 2) >> 2;
      gp[i] = gc[i];
      gc[i] = gn[i];
    }
    emit(2 * y, top);
    emit(2 * y + 1, bot);
  }
}

// Fast path (w % 4 == 0, word-aligned frames): one thread = 4 source pixels (3 words + one halo word each side) x RZ_ROWS
// source rows -> 8 output pixels (24 bytes = three 8-byte stores) x 2*RZ_ROWS rows.  All arithmetic runs on two 16-bit
// lanes per register (g <= 1020, 3g <= 3060: no carries between lanes), bytes are moved with PRMT only:
//   e[-4..15] = the 5 words; A/B = byte pairs (e[i], e[i+1]) at even / odd i; gE = 3A + B(i-3), gO = 3A + B(i+3)
//   G[j] = the g pair of output bytes (2j, 2j+1); vertical pass per G register; one PRMT packs four results.
__host__ __device__ constexpr int rz_src(int ob) { return 3 * ((ob / 3) >> 1) + ob % 3; }   // near source byte of output byte
__host__ __device__ constexpr int rz_odd(int ob) { return (ob / 3) & 1; }                    // output pixel parity

__device__ __forceinline__ uint32_t rz_pair16(const uint32_t (&X)[5], int i) {               // e[i] | e[i+1] << 16
  const int p = i + 4, wa = p >> 2, ba = p & 3;
  if (ba < 3) return __byte_perm(X[wa], 0u, (uint32_t)(ba | 0x40 | ((ba + 1) << 8) | 0x4000));
  return __byte_perm(X[wa], X[wa + 1], 0x0403u) & 0x00FF00FFu;
}

template <bool LINEAR>
__global__ void __launch_bounds__(RZ_THREADS) resize2x_px4_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                                  int h, int w, int rows, bool wide_ok) {
  const int units = w >> 2;
  const int u = blockIdx.x * RZ_THREADS + threadIdx.x;
  if (u >= units) return;
  const int row_in = 3 * w, row_out = 6 * w;
  const int y0 = blockIdx.y * rows, y1 = min(h, y0 + rows);
  const uint8_t* src = in + (size_t)blockIdx.z * h * row_in + 12 * u;
  uint8_t* dst = out + (size_t)blockIdx.z * 2 * h * row_out + 24 * u;
  const bool first = (u == 0), last = (u == units - 1);

  auto load_words = [&](int y, uint32_t (&X)[5]) {
    const uint32_t* r = reinterpret_cast<const uint32_t*>(src + (size_t)y * row_in);
    X[1] = __ldg(r);
    X[2] = __ldg(r + 1);
    X[3] = __ldg(r + 2);
    if (LINEAR) {
      X[0] = first ? (X[1] << 8) : __ldg(r - 1);      // e[-3..-1]: left neighbour pixel, replicated at the border
      X[4] = last ? (X[3] >> 8) : __ldg(r + 3);       // e[12..14]: right neighbour pixel
    } else {
      X[0] = 0u;
      X[4] = 0u;
    }
  };
  // A warp's 32 x 24 output bytes of one row are contiguous: full warps stage them in shared memory and store 16 bytes per
  // lane, fully coalesced (48 x 16 B); 8-byte stores at a 24-byte lane stride are the fall-back (partial warps, w % 8 != 0).
  __shared__ __align__(16) uint8_t stage_all[RZ_THREADS / 32][768];
  const int lane = threadIdx.x & 31;
  uint8_t* stage = stage_all[threadIdx.x >> 5];
  // LINEAR keeps the direct stores: staging cost it 16 registers (64 -> 80, 8 -> 6 blocks per SM) and 20 % (measured)
  const bool staged = !LINEAR && wide_ok && (u - lane + 32 <= units);
  constexpr int copies = LINEAR ? 1 : 2;            // replication writes every row twice
  auto store_row = [&](int dy, const uint32_t (&o)[6]) {
    if (staged) {
      uint2* sp = reinterpret_cast<uint2*>(stage + 24 * lane);
      sp[0] = make_uint2(o[0], o[1]);
      sp[1] = make_uint2(o[2], o[3]);
      sp[2] = make_uint2(o[4], o[5]);
      __syncwarp();
      const uint4* s4 = reinterpret_cast<const uint4*>(stage);
      const uint4 v0 = s4[lane], v1 = s4[32 + (lane & 15)];
      __syncwarp();
#pragma unroll
      for (int c = 0; c < copies; ++c) {
        uint4* d4 = reinterpret_cast<uint4*>(dst - 24 * lane + (size_t)(dy + c) * row_out);
        d4[lane] = v0;
        if (lane < 16) d4[32 + lane] = v1;
      }
    } else {
#pragma unroll
      for (int c = 0; c < copies; ++c) {
        uint2* d = reinterpret_cast<uint2*>(dst + (size_t)(dy + c) * row_out);
        d[0] = make_uint2(o[0], o[1]);
        d[1] = make_uint2(o[2], o[3]);
        d[2] = make_uint2(o[4], o[5]);
      }
    }
  };

  if (!LINEAR) {
    for (int y = y0; y < y1; ++y) {
      uint32_t X[5], o[6];
      load_words(y, X);
#pragma unroll
      for (int q = 0; q < 6; ++q) {
        int wb = 5;                                     // lowest word any of the four source bytes lies in (they span <= 2)
#pragma unroll
        for (int t = 0; t < 4; ++t) wb = min(wb, (rz_src(4 * q + t) + 4) >> 2);
        uint32_t sel = 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int p = rz_src(4 * q + t) + 4;
          sel |= (uint32_t)(((p >> 2) - wb) * 4 + (p & 3)) << (4 * t);
        }
        o[q] = __byte_perm(X[wb], X[wb + 1], sel);
      }
      store_row(2 * y, o);
    }
    return;
  }

  auto load_g = [&](int y, uint32_t (&G)[12]) {
    uint32_t X[5], A[6], B[9], gE[6], gO[6];
    load_words(y, X);
#pragma unroll
    for (int p = 0; p < 6; ++p) A[p] = rz_pair16(X, 2 * p);
#pragma unroll
    for (int m = 0; m < 9; ++m) B[m] = rz_pair16(X, 2 * m - 3);
#pragma unroll
    for (int p = 0; p < 6; ++p) {
      gE[p] = 3u * A[p] + B[p];
      gO[p] = 3u * A[p] + B[p + 3];
    }
#pragma unroll
    for (int j = 0; j < 12; ++j) {
      const int i0 = rz_src(2 * j), i1 = rz_src(2 * j + 1);
      const uint32_t ra = rz_odd(2 * j) ? gO[i0 >> 1] : gE[i0 >> 1];
      const uint32_t rb = rz_odd(2 * j + 1) ? gO[i1 >> 1] : gE[i1 >> 1];
      const uint32_t la = 2 * (i0 & 1), lb = 4 + 2 * (i1 & 1);
      G[j] = __byte_perm(ra, rb, la | ((la + 1) << 4) | (lb << 8) | ((lb + 1) << 12));
    }
  };

  // Per output pair j the state between rows is q(y-1), q(y) (alternating in Qa / Qb) and m(y), with q = (g >> 2) and
  // m = (3g >> 2) computed once per source row.  The low lane of m carries two stray bits (14, 15) from the high lane;
  // q is masked, so the sum cannot carry out of the lane, and PRMT below only takes bits 0..7 of each lane.
  constexpr uint32_t MASK = 0x3FFF3FFFu, HALF = 0x00020002u;
  uint32_t Qa[12], Qb[12], M[12], G[12];
  load_g(max(y0 - 1, 0), G);
#pragma unroll
  for (int j = 0; j < 12; ++j) Qa[j] = (G[j] >> 2) & MASK;
  load_g(y0, G);
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    Qb[j] = (G[j] >> 2) & MASK;
    M[j] = (3u * G[j]) >> 2;
  }
  auto step = [&](int y, uint32_t (&qp)[12]) {      // in: qp = q(y-1), M = m(y); out: qp = q(y+1), M = m(y+1)
    load_g(min(y + 1, h - 1), G);
    uint32_t top[12], bot[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) {
      const uint32_t qn = (G[j] >> 2) & MASK;
      top[j] = (M[j] + qp[j] + HALF) >> 2;
      bot[j] = (M[j] + qn + HALF) >> 2;
      qp[j] = qn;
      M[j] = (3u * G[j]) >> 2;
    }
    uint32_t o[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) o[q] = __byte_perm(top[2 * q], top[2 * q + 1], 0x6420u);
    store_row(2 * y, o);
#pragma unroll
    for (int q = 0; q < 6; ++q) o[q] = __byte_perm(bot[2 * q], bot[2 * q + 1], 0x6420u);
    store_row(2 * y + 1, o);
  };
  for (int y = y0; y < y1; y += 2) {
    step(y, Qa);
    if (y + 1 < y1) step(y + 1, Qb);
  }
}

int launch_resize2x(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, int linear, cudaStream_t s) {
  if (!in || !out || n <= 0 || h <= 0 || w <= 0) return fail(ctx, VEX_ERR_INVALID, "vex_resize2x_u8: bad argument");
  const int words = (6 * w + 3) / 4;
  dim3 grid((words + RZ_THREADS - 1) / RZ_THREADS, (h + RZ_ROWS - 1) / RZ_ROWS, n);
  if (grid.y > 65535 || grid.z > 65535) return fail(ctx, VEX_ERR_INVALID, "vex_resize2x_u8: frame or batch too large");
  // 6w bytes per output row: rows stay word-aligned iff w is even (cudaMalloc'd frames; checked for sub-buffers)
  const bool word_ok = (w % 2 == 0) && (reinterpret_cast<uintptr_t>(out) % 4 == 0);
  if (w % 4 == 0 && reinterpret_cast<uintptr_t>(in) % 4 == 0 && reinterpret_cast<uintptr_t>(out) % 8 == 0 &&
      !ctx->opt_force_generic_resize) {
    // the H pass of the rows above and below a strip is recomputed (2 extra rows per strip): 16-row strips when there are
    // enough threads to fill the SMs several times over, 8-row strips otherwise
    const int rows = ((long long)n * (h / 16 + 1) * (w / 4) >= 64LL * 2048 * ctx->sm_count / 148) ? 16 : RZ_ROWS;
    dim3 g4((w / 4 + RZ_THREADS - 1) / RZ_THREADS, (h + rows - 1) / rows, n);
    const bool wide_ok = (w % 8 == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0);   // 16-byte aligned output rows
    if (linear) resize2x_px4_kernel<true><<<g4, RZ_THREADS, 0, s>>>(in, out, h, w, rows, wide_ok);
    else resize2x_px4_kernel<false><<<g4, RZ_THREADS, 0, s>>>(in, out, h, w, rows, wide_ok);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    return VEX_OK;
  }
  if (linear) {
    if (word_ok) resize2x_kernel<true, true><<<grid, RZ_THREADS, 0, s>>>(in, out, h, w);
    else resize2x_kernel<true, false><<<grid, RZ_THREADS, 0, s>>>(in, out, h, w);
  } else {
    if (word_ok) resize2x_kernel<false, true><<<grid, RZ_THREADS, 0, s>>>(in, out, h, w);
    else resize2x_kernel<false, false><<<grid, RZ_THREADS, 0, s>>>(in, out, h, w);
  }
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  return VEX_OK;
}

}  // namespace vex
