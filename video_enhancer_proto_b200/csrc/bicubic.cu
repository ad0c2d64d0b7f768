// 2x bicubic upscale, uint8 HWC -> uint8 HWC, for sm_100a.  HBM-bound: 3*H*W bytes in, 12*H*W out.
//
// Replaces GPUVideoEnhancer._enhance_frame_cpu_fallback (reference app.py:359-385):
//   F.interpolate(x/255, scale_factor=2, mode='bicubic', align_corners=False) -> clip(*255) -> uint8
// and its OpenCV twin cv2.resize(.., INTER_CUBIC) (basic_upscaler_fallback.py:142).
//
// Math.  out index d -> source s=(d+0.5)/2-0.5, i=floor(s), t=s-i in {0.75 (d even), 0.25 (d odd)};
// taps i-1..i+2 clamped to the image (border replicate); Keys cubic with A=-0.75 gives the exact
// dyadic weights  t=0.75: [-9, 67, 225, -27]/256   t=0.25: [-27, 225, 67, -9]/256.
// Separable (x then y).  All products are carried as integers (weights*256) in fp32: the horizontal
// sum is < 2^17 and every in-range vertical sum is < 2^24, so fp32 FMAs are EXACT; the result is the
// mathematically exact value of the interpolant, scaled by 65536.  round_mode 0 truncates
// (app.py:376 astype(uint8) after clip), round_mode 1 rounds half up (cv2's fixed-point rounding).
// The reference's float path differs from the exact value only by fp32 rounding at integer
// boundaries (<= 1 LSB, see tests/test_bicubic_oracle.py).
#include "vex_internal.cuh"

namespace vex {
namespace {

constexpr int TPX = 128;       // input pixels per tile in x
constexpr int TR = 32;         // input rows per tile
constexpr int RG = 8;          // input rows per row group (one warp pair)
constexpr int NTHREADS = 256;  // 64 column groups (2 input px each) x 4 row groups
constexpr int IN_VECS = (TPX * 3 + 32) / 16;   // 26 x 16B per smem row: [X0*3-16, X0*3+400)
constexpr int PITCH = IN_VECS * 16;            // 416
constexpr int SROWS = TR + 4;                  // rows R0-2 .. R0+TR+1
constexpr int STAGE_BYTES = 2 * 384;           // per warp: 2 output rows x 64 in-px x 6 B

__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_v4(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ float byte_f(uint32_t w, int k) { return (float)((w >> (8 * k)) & 0xffu); }

// horizontal pass for one smem row: 18 bytes (6 px) -> 12 values (4 out px x 3 ch), scaled by 256
__device__ __forceinline__ void hpass(const uint8_t* srow, int wi, int sh, float (&h)[12]) {
  const uint32_t* wp = reinterpret_cast<const uint32_t*>(srow) + wi;
  uint32_t w[6];
#pragma unroll
  for (int j = 0; j < 5; ++j) w[j] = wp[j];
  w[5] = 0;
#pragma unroll
  for (int j = 0; j < 5; ++j) w[j] = __funnelshift_r(w[j], w[j + 1], sh);
  float p[18];
#pragma unroll
  for (int k = 0; k < 18; ++k) p[k] = byte_f(w[k >> 2], k & 3);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // even out px (t=0.75) of input px q0+2 and q0+3 ; odd out px (t=0.25)
    h[0 + c] = -9.f * p[0 + c] + 67.f * p[3 + c] + 225.f * p[6 + c] - 27.f * p[9 + c];
    h[3 + c] = -27.f * p[3 + c] + 225.f * p[6 + c] + 67.f * p[9 + c] - 9.f * p[12 + c];
    h[6 + c] = -9.f * p[3 + c] + 67.f * p[6 + c] + 225.f * p[9 + c] - 27.f * p[12 + c];
    h[9 + c] = -27.f * p[6 + c] + 225.f * p[9 + c] + 67.f * p[12 + c] - 9.f * p[15 + c];
  }
}

template <int ROUND>
__device__ __forceinline__ uint32_t quant(float v) {
  // v = exact value * 65536.  clip to [0,255] then truncate (ROUND=0) or round half up (ROUND=1).
  float s = fmaf(v, 1.0f / 65536.0f, ROUND ? 0.5f : 0.0f);
  return min(__float2uint_rz(s), 255u);   // float->uint conversion clamps negatives to 0
}

template <int ROUND>
__device__ __forceinline__ void emit_row(uint8_t* stage_row, int lane, const float (&a)[12],
                                         const float (&b)[12], const float (&c)[12],
                                         const float (&d)[12], float w0, float w1, float w2,
                                         float w3) {
  uint32_t o[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int e = 4 * j + k;
      // negatives first keeps every partial sum of an in-range result below 2^24 (exact)
      float v = w0 * a[e] + w3 * d[e];
      v = (w1 < w2) ? fmaf(w1, b[e], v) : fmaf(w2, c[e], v);
      v = (w1 < w2) ? fmaf(w2, c[e], v) : fmaf(w1, b[e], v);
      word |= quant<ROUND>(v) << (8 * k);
    }
    o[j] = word;
  }
  uint32_t* sp = reinterpret_cast<uint32_t*>(stage_row) + 3 * lane;
  sp[0] = o[0];
  sp[1] = o[1];
  sp[2] = o[2];
}

// Fast path: W % 16 == 0 (so 3W and 6W are multiples of 16) and 16B-aligned buffers.
template <int ROUND>
__global__ void __launch_bounds__(NTHREADS, 2)
bicubic2x_tile_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W) {
  __shared__ __align__(16) uint8_t s_in[SROWS * PITCH];
  __shared__ __align__(16) uint8_t s_stage[(NTHREADS / 32) * STAGE_BYTES];

  const int tid = threadIdx.x;
  const int X0 = blockIdx.x * TPX;
  const int R0 = blockIdx.y * TR;
  const size_t in_row_bytes = (size_t)W * 3;
  const size_t out_row_bytes = (size_t)W * 6;
  const uint8_t* fin = in + (size_t)blockIdx.z * H * in_row_bytes;
  uint8_t* fout = out + (size_t)blockIdx.z * (2 * (size_t)H) * out_row_bytes;

  // ---- stage the input tile (+2 px / +2 rows halo) with coalesced 16B loads
  const int cb0 = X0 * 3 - 16;
  for (int idx = tid; idx < SROWS * IN_VECS; idx += NTHREADS) {
    const int rr = idx / IN_VECS, v = idx - rr * IN_VECS;
    int y = R0 - 2 + rr;
    y = min(max(y, 0), H - 1);
    const int cb = cb0 + 16 * v;
    uint4 val = make_uint4(0, 0, 0, 0);
    if (cb >= 0 && cb + 16 <= (int)in_row_bytes) val = ldg_nc_v4(fin + (size_t)y * in_row_bytes + cb);
    *reinterpret_cast<uint4*>(s_in + rr * PITCH + 16 * v) = val;
  }
  __syncthreads();
  // ---- border replicate in x (only tiles touching the left/right image edge)
  const bool left = (X0 == 0), right = (X0 + TPX + 2 > W);
  if (left || right) {
    if (tid < SROWS) {
      uint8_t* r = s_in + tid * PITCH;
      if (left) {
#pragma unroll
        for (int c = 0; c < 3; ++c) r[10 + c] = r[13 + c] = r[16 + c];
      }
      if (right) {
        const int sb = 16 + (W - 1 - X0) * 3;   // last real pixel
#pragma unroll
        for (int c = 0; c < 3; ++c) r[sb + 3 + c] = r[sb + 6 + c] = r[sb + c];
      }
    }
    __syncthreads();
  }

  // ---- compute: thread = 2 input px (4 out px = 12 B) x RG input rows (2*RG out rows)
  const int lane = tid & 31, warp = tid >> 5;
  const int tx = tid & 63, rg = tid >> 6;
  const int sbase = 10 + 6 * tx;
  const int wi = sbase >> 2, sh = (sbase & 3) * 8;
  uint8_t* stage = s_stage + warp * STAGE_BYTES;
  const int Xw = X0 + 64 * (warp & 1);                 // first input px of this warp
  const int valid_vecs = min(24, max(0, (W - Xw) * 6 / 16));

  float h0[12], h1[12], h2[12], h3[12], h4[12];
  const uint8_t* srow = s_in + (rg * RG) * PITCH;
  hpass(srow + 0 * PITCH, wi, sh, h0);
  hpass(srow + 1 * PITCH, wi, sh, h1);
  hpass(srow + 2 * PITCH, wi, sh, h2);
  hpass(srow + 3 * PITCH, wi, sh, h3);
#pragma unroll
  for (int j = 0; j < RG; ++j) {
    hpass(srow + (j + 4) * PITCH, wi, sh, h4);
    // out row 2i (t=0.75): rows i-2..i+1 ; out row 2i+1 (t=0.25): rows i-1..i+2
    emit_row<ROUND>(stage, lane, h0, h1, h2, h3, -9.f, 67.f, 225.f, -27.f);
    emit_row<ROUND>(stage + 384, lane, h1, h2, h3, h4, -27.f, 225.f, 67.f, -9.f);
    __syncwarp();
    const int i = R0 + rg * RG + j;
    if (i < H) {
      uint8_t* orow = fout + (size_t)(2 * i) * out_row_bytes + (size_t)Xw * 6;
      // 48 vectors (2 rows x 24): lanes 0..31 then 0..15
      {
        const int v = lane;                       // 0..31
        const int row = v >= 24, vv = row ? v - 24 : v;
        if (vv < valid_vecs)
          stg_v4(orow + row * out_row_bytes + 16 * vv,
                 *reinterpret_cast<const uint4*>(stage + row * 384 + 16 * vv));
      }
      if (lane < 16) {
        const int vv = lane + 8;
        if (vv < valid_vecs)
          stg_v4(orow + out_row_bytes + 16 * vv,
                 *reinterpret_cast<const uint4*>(stage + 384 + 16 * vv));
      }
    }
    __syncwarp();
#pragma unroll
    for (int e = 0; e < 12; ++e) {
      h0[e] = h1[e];
      h1[e] = h2[e];
      h2[e] = h3[e];
      h3[e] = h4[e];
    }
  }
}

// Generic path: any H, W >= 1 and any alignment.  One thread per output byte.
template <int ROUND>
__global__ void __launch_bounds__(256)
bicubic2x_generic_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int n, int H,
                         int W) {
  const size_t total = (size_t)n * (2 * (size_t)H) * (2 * (size_t)W) * 3;
  const float W075[4] = {-9.f, 67.f, 225.f, -27.f};
  const float W025[4] = {-27.f, 225.f, 67.f, -9.f};
  for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < total;
       j += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(j % 3);
    size_t t = j / 3;
    const int dx = (int)(t % (2 * (size_t)W));
    t /= (2 * (size_t)W);
    const int dy = (int)(t % (2 * (size_t)H));
    const size_t f = t / (2 * (size_t)H);
    const int ix = (dx & 1) ? (dx >> 1) : (dx >> 1) - 1;   // floor(s)
    const int iy = (dy & 1) ? (dy >> 1) : (dy >> 1) - 1;
    const float* wx = (dx & 1) ? W025 : W075;
    const float* wy = (dy & 1) ? W025 : W075;
    const uint8_t* fin = in + f * (size_t)H * W * 3;
    float hv[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int y = min(max(iy - 1 + a, 0), H - 1);
      float acc = 0.f;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int x = min(max(ix - 1 + b, 0), W - 1);
        acc = fmaf(wx[b], (float)fin[((size_t)y * W + x) * 3 + c], acc);
      }
      hv[a] = acc;
    }
    // same exactness argument as the tile kernel: negative taps first
    float v;
    if (dy & 1) { v = wy[0] * hv[0] + wy[3] * hv[3]; v = fmaf(wy[2], hv[2], v); v = fmaf(wy[1], hv[1], v); }
    else        { v = wy[0] * hv[0] + wy[3] * hv[3]; v = fmaf(wy[1], hv[1], v); v = fmaf(wy[2], hv[2], v); }
    out[j] = (uint8_t)quant<ROUND>(v);
  }
}

}  // namespace

int launch_bicubic2x(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w,
                     int round_mode, cudaStream_t s) {
  if (!in || !out || n <= 0 || h <= 0 || w <= 0 || (round_mode != 0 && round_mode != 1))
    return fail(ctx, VEX_ERR_INVALID, "vex_bicubic2x_u8: bad argument");
  const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  if (w % 16 == 0 && aligned && !ctx->opt_force_generic_bicubic) {
    for (int f0 = 0; f0 < n; f0 += 65535) {
      const int nf = min(n - f0, 65535);
      dim3 grid((w + TPX - 1) / TPX, (h + TR - 1) / TR, nf);
      const uint8_t* pin = in + (size_t)f0 * h * w * 3;
      uint8_t* pout = out + (size_t)f0 * h * w * 12;
      if (round_mode) bicubic2x_tile_kernel<1><<<grid, NTHREADS, 0, s>>>(pin, pout, h, w);
      else            bicubic2x_tile_kernel<0><<<grid, NTHREADS, 0, s>>>(pin, pout, h, w);
      ctx->launches++;
    }
  } else {
    const size_t total = (size_t)n * h * w * 12;
    const int blocks = (int)min((total + 255) / 256, (size_t)ctx->sm_count * 32);
    if (round_mode) bicubic2x_generic_kernel<1><<<blocks, 256, 0, s>>>(in, out, n, h, w);
    else            bicubic2x_generic_kernel<0><<<blocks, 256, 0, s>>>(in, out, n, h, w);
    ctx->launches++;
  }
  VEX_CUDA(ctx, cudaGetLastError());
  return VEX_OK;
}

}  // namespace vex
