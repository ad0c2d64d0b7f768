// 2x bicubic upscale, uint8 HWC -> uint8 HWC, for sm_100a.  Algorithmic HBM traffic: 3*H*W bytes in, 12*H*W out.
//
// Replaces GPUVideoEnhancer._enhance_frame_cpu_fallback (reference app.py:359-385):
//   F.interpolate(x/255, scale_factor=2, mode='bicubic', align_corners=False) -> clip(*255) -> uint8
// and its OpenCV twin cv2.resize(.., INTER_CUBIC) (basic_upscaler_fallback.py:142).
//
// Math.  out index d -> source s=(d+0.5)/2-0.5, i=floor(s), t=s-i in {0.75 (d even), 0.25 (d odd)};
// taps i-1..i+2 clamped to the image (border replicate); Keys cubic with A=-0.75 gives the exact
// dyadic weights  t=0.75: [-9, 67, 225, -27]/256   t=0.25: [-27, 225, 67, -9]/256.
// Separable (x then y), carried as integers in fp32: the horizontal sums are integers < 2^17, the
// vertical sums are multiples of 2^-16 below 2^8, so every FMA is exact as long as the running sum
// stays below 2^24 units.  The vertical chain adds the 225-tap first, then the two negative taps,
// then the 67-tap, so for ordinary content the running sum never exceeds the final value; it can
// leave the exact range only when a near-saturated 2-px stripe row sits next to rows of the inverse
// stripe (negative h on a negative-weight tap), and then the result is perturbed by 2^-16 LSB,
// which is visible only on an exact tie.  round_mode 0 truncates (app.py:376 astype(uint8) after clip),
// round_mode 1 rounds half up (cv2's fixed-point rounding).  The reference's float path differs from
// the exact value only by fp32 rounding at integer boundaries (<= 1 LSB, tests/test_bicubic_gpu.py).
//
// Why it looks like this.  With 16 taps per output byte and 4 output bytes per input byte the kernel is
// CUDA-core issue-bound long before it is HBM-bound (measured on B200, tools/ubench/pipes.cu: FFMA ~90,
// FFMA2 63 x 2 lanes, ALU-pipe ops 63, F2I 16 (!), F2IP.U8.F32 66 x 2 lanes per clock per SM; at 80 % of HBM
// an SM must emit ~15 output bytes per clock, i.e. < 9 issue slots per byte).  So:
//   * all arithmetic is packed fp32x2 (FFMA2): one thread runs two column strips in the two lanes;
//   * uint8 -> fp32 is ONE byte-permute: PRMT builds the float 2^15 + p, and the bias (256 * 2^15 = 2^23)
//     is removed exactly by the addend of the first FMA of the horizontal chain (weights sum to 256);
//   * the even and the odd output of a tap window share the 225-tap product: 7 FMAs per output pair;
//   * fp32 -> uint8 with clamp and pack is F2IP.U8.F32.TRUNC: two values per instruction
//     (cvt.rzi.s32.f32 + cvt.pack.sat.u8.s32 fuse in ptxas), rounding folded into the horizontal bias;
//   * a thread walks down a strip of rows keeping a 4-row window of horizontal sums in registers
//     (row halo cost 4/R), input rows come straight from global/L1 one row ahead, output rows are
//     staged per warp in shared memory and leave as 16-byte fully coalesced stores (768 B per warp row).
#include "vex_internal.cuh"

#include <algorithm>

namespace vex {
namespace {

typedef unsigned long long u64;

__device__ __forceinline__ u64 pk(uint32_t lo, uint32_t hi) {
  u64 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void unpk(u64 v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 bc(float w) {   // broadcast constant to both lanes (ptxas folds it to an immediate)
  const uint32_t u = __float_as_uint(w);
  return pk(u, u);
}
__device__ __forceinline__ u64 fma2(u64 a, float w, u64 c) {
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(bc(w)), "l"(c));
  return d;
}
__device__ __forceinline__ u64 mul2(u64 a, float w) {
  u64 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(bc(w)));
  return d;
}
// four floats (already in [0,256) scale) -> 4 clamped, truncated bytes, b0 at the lowest address.
// cvt.pack: d = (c << 16) | sat(a) << 8 | sat(b)   (measured, tools/ubench/packtest.cu)
__device__ __forceinline__ uint32_t pack4(float b0, float b1, float b2, float b3) {
  const int i0 = __float2int_rz(b0), i1 = __float2int_rz(b1), i2 = __float2int_rz(b2), i3 = __float2int_rz(b3);
  uint32_t hi, r;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, 0;" : "=r"(hi) : "r"(i3), "r"(i2));
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(i1), "r"(i0), "r"(hi));
  return r;
}

constexpr int STRIP_WARPS = 1;            // warps per block: one, so that every loop bound and ring slot is provably
                                          // warp-uniform (uniform datapath, no divergence check around __syncwarp)
constexpr int RESIDENT_WARPS = 12;        // per SM at 168 registers
constexpr int WARP_PX = 128;              // input pixels per warp row segment: lane = 4 px = 2 strips x 2 px
constexpr int STAGE_ROW = WARP_PX * 6;    // 768 output bytes per warp per output row
constexpr int RING_D = 6;                 // input rows in flight per warp (cp.async groups)
constexpr int RING_SLOTS = 8;             // ring slots (power of two > RING_D; slot = step & 7)
constexpr int RING_ROW = 416;             // 26 x 16 B: bytes [S-16, S+400) of a row
constexpr float VS = 1.0f / 65536.0f;     // vertical weights carry the 2^-16 so the sum lands in [0,256)

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
  return d;
}
// byte k of w -> the float 2^15 + byte (byte lands in mantissa bits 8..15 of 0x47000000)
// ZSEL picks which zero byte of the magic constant fills byte 0 / 2 of the result: a byte that feeds both strips is
// converted once per strip with different (equivalent) selectors on purpose - with identical selectors ptxas shares
// the PRMT and then pays several MOVs per shared value to build the two aligned register pairs.
template <int IDX, int ZSEL>
__device__ __forceinline__ uint32_t magic_f(const uint32_t (&raw)[7]) {
  return prmt(raw[IDX >> 2], 0x47000000u, 0x7000u | (ZSEL << 8) | ((IDX & 3) << 4) | ZSEL);
}

// this lane's destinations of an output row pair (2 rows x 48 vectors of 16 B per warp): vector `lane` of each row,
// and lanes 0..15 also vector 32 + lane.  Two running pointers and immediate offsets - no per-lane offset registers.
struct LaneOut {
  uint8_t* pa;          // running: output row 2i, vector `lane`
  uint8_t* pb;          // running: output row 2i + 1, vector `lane`
  bool ok1, ok2;        // vector `lane` / vector 32 + lane (lanes 0..15) lies inside the row
};
__device__ __forceinline__ void stg_v4_if(void* p, const uint4& v, bool pred) {
  asm volatile(
      "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};\n\t}"
      ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"((int)pred)
      : "memory");
}

// One input row of one warp segment.  raw[] = the 28 bytes [12*lane-8, 12*lane+20) of the row segment:
// pixels 4*lane-2 .. 4*lane+5 start at byte index 2.  Lane = strip A (px 4l, 4l+1; fp32x2 low lane) +
// strip B (px 4l+2, 4l+3; high lane).  Produces the 12 horizontal sums per strip (out px 2u+e, channel c
// at j = (2u+e)*3+c), and, if EMIT, output rows 2i and 2i+1 of input row i whose window is
// S[(K+2)&3] = row i-2, S[(K+3)&3] = i-1, S[K] = i, S[(K+1)&3] = i+1, new = i+2.
template <int ROUND, int K, bool EMIT>
__device__ __forceinline__ void strip_step(u64 (&S)[4][12], uint32_t (&raw)[7], bool left, bool right,
                                           uint8_t* stage, const LaneOut& o, bool store_ok, int lane) {
  // border replicate in x on the raw bytes (app.py:372 -> ATen clamps tap indices)
  if (left) {            // px -2, -1 <- px 0 : bytes 2..7 <- bytes 8,9,10
    raw[0] = prmt(raw[2], raw[2], 0x1000u);
    raw[1] = prmt(raw[2], raw[2], 0x2102u);
  }
  if (right) {           // px +4, +5 <- px 3 : bytes 20..25 <- bytes 17,18,19
    raw[5] = prmt(raw[4], raw[4], 0x1321u);
    raw[6] = prmt(raw[4], raw[4], 0x0032u);
  }
  u64 P[6][3];
#define VEX_CONV(T, C) P[T][C] = pk(magic_f<3 * T + C + 2, 4>(raw), magic_f<3 * T + C + 8, 5>(raw));
#define VEX_CONV3(T) VEX_CONV(T, 0) VEX_CONV(T, 1) VEX_CONV(T, 2)
  VEX_CONV3(0) VEX_CONV3(1) VEX_CONV3(2) VEX_CONV3(3) VEX_CONV3(4) VEX_CONV3(5)
#undef VEX_CONV3
#undef VEX_CONV
  // sum of weights = 256, so the 2^15 conversion bias contributes exactly 2^23; round-half-up = +128 here
  const u64 bias = bc(-8388608.0f + (ROUND ? 128.0f : 0.0f));
  constexpr int SA = (K + 2) & 3, SB = (K + 3) & 3, SC = K, SD = (K + 1) & 3;
  float ea[12], eb[12], oa[12], ob[12];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int j1 = 6 * u + c, j2 = j1 + 3;     // even / odd output pixel of input pixel u
      u64 X1, X2, T1, T2;
      if constexpr (EMIT) {
        // start both vertical chains with the row that is about to be overwritten (row i-2)
        X1 = mul2(S[SC][j1], 225.f * VS);
        X2 = mul2(S[SC][j2], 225.f * VS);
        T1 = fma2(S[SA][j1], -9.f * VS, X1);
        T2 = fma2(S[SA][j2], -9.f * VS, X2);
      }
      const u64 Xh = fma2(P[u + 2][c], 225.f, bias);
      const u64 hE = fma2(P[u + 1][c], 67.f, fma2(P[u + 3][c], -27.f, fma2(P[u][c], -9.f, Xh)));
      const u64 hO = fma2(P[u + 3][c], 67.f, fma2(P[u + 4][c], -9.f, fma2(P[u + 1][c], -27.f, Xh)));
      S[SA][j1] = hE;
      S[SA][j2] = hO;
      if constexpr (EMIT) {
        const u64 E1 = fma2(S[SB][j1], 67.f * VS, fma2(S[SD][j1], -27.f * VS, T1));
        const u64 E2 = fma2(S[SB][j2], 67.f * VS, fma2(S[SD][j2], -27.f * VS, T2));
        const u64 O1 = fma2(S[SD][j1], 67.f * VS, fma2(hE, -9.f * VS, fma2(S[SB][j1], -27.f * VS, X1)));
        const u64 O2 = fma2(S[SD][j2], 67.f * VS, fma2(hO, -9.f * VS, fma2(S[SB][j2], -27.f * VS, X2)));
        unpk(E1, ea[j1], eb[j1]);
        unpk(E2, ea[j2], eb[j2]);
        unpk(O1, oa[j1], ob[j1]);
        unpk(O2, oa[j2], ob[j2]);
      }
    }
  }
  if constexpr (EMIT) {
    // lane's 24 output bytes per row: strip A (12 B) then strip B (12 B)
    uint2* se = reinterpret_cast<uint2*>(stage + 24 * lane);
    uint2* so = reinterpret_cast<uint2*>(stage + STAGE_ROW + 24 * lane);
    {
      const uint32_t a0 = pack4(ea[0], ea[1], ea[2], ea[3]), a1 = pack4(ea[4], ea[5], ea[6], ea[7]),
                     a2 = pack4(ea[8], ea[9], ea[10], ea[11]);
      const uint32_t b0 = pack4(eb[0], eb[1], eb[2], eb[3]), b1 = pack4(eb[4], eb[5], eb[6], eb[7]),
                     b2 = pack4(eb[8], eb[9], eb[10], eb[11]);
      se[0] = make_uint2(a0, a1);
      se[1] = make_uint2(a2, b0);
      se[2] = make_uint2(b1, b2);
    }
    {
      const uint32_t a0 = pack4(oa[0], oa[1], oa[2], oa[3]), a1 = pack4(oa[4], oa[5], oa[6], oa[7]),
                     a2 = pack4(oa[8], oa[9], oa[10], oa[11]);
      const uint32_t b0 = pack4(ob[0], ob[1], ob[2], ob[3]), b1 = pack4(ob[4], ob[5], ob[6], ob[7]),
                     b2 = pack4(ob[8], ob[9], ob[10], ob[11]);
      so[0] = make_uint2(a0, a1);
      so[1] = make_uint2(a2, b0);
      so[2] = make_uint2(b1, b2);
    }
    __syncwarp();
    // 96 x 16 B (2 rows x 48) leave in four coalesced passes (two of them half warps)
    const uint4* sv = reinterpret_cast<const uint4*>(stage);
    const bool k1 = o.ok1 && store_ok, k2 = o.ok2 && store_ok;
    stg_v4_if(o.pa, sv[lane], k1);
    stg_v4_if(o.pb, sv[48 + lane], k1);
    stg_v4_if(o.pa + 512, sv[32 + lane], k2);
    stg_v4_if(o.pb + 512, sv[80 + lane], k2);
  }
}

// Fast path: W % 16 == 0 (3W and 6W multiples of 16) and 16B-aligned buffers.  R = rows per strip (multiple of 4).
// Input rows reach the lanes through a per-warp cp.async ring in shared memory (RING_D rows in flight, no
// registers held while they fly): 26 lanes copy the warp's 416-byte row segment [S-16, S+400) as 16-byte chunks,
// chunks outside the row are zero-filled (src-size 0) and rebuilt by the border replicate.
template <int ROUND>
__global__ void __launch_bounds__(STRIP_WARPS * 32, RESIDENT_WARPS / STRIP_WARPS)
bicubic2x_strip_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W, int R,
                       int wpr, int nstrips, int total_warps) {
  __shared__ __align__(16) uint8_t s_stage[STRIP_WARPS][2 * STAGE_ROW];
  __shared__ __align__(16) uint8_t s_ring[STRIP_WARPS][RING_SLOTS][RING_ROW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gw = blockIdx.x * STRIP_WARPS + warp;
  if (gw >= total_warps) return;
  const int xw = gw % wpr;
  const int t = gw / wpr;
  const int strip = t % nstrips, f = t / nstrips;
  const int r0 = strip * R;
  const int px0 = xw * WARP_PX + 4 * lane;
  const bool left = px0 == 0, right = px0 + 4 == W;
  const uint32_t in_row = (uint32_t)W * 3;
  const size_t out_row = (size_t)W * 6;
  const int valid_vecs = min(48, (W - xw * WARP_PX) * 6 / 16);

  // cp.async source of this lane: chunk `lane` of the segment, if it lies inside the row
  const int chunk_off = xw * (WARP_PX * 3) - 16 + 16 * lane;
  const bool chunk_lane = lane < RING_ROW / 16;
  const uint32_t chunk_bytes = (chunk_lane && chunk_off >= 0 && chunk_off < (int)in_row) ? 16u : 0u;
  const uint8_t* fsrc = in + (size_t)f * H * in_row + (chunk_bytes ? chunk_off : 0);
  const uint32_t ring_base = (uint32_t)__cvta_generic_to_shared(&s_ring[warp][0][0]);
  const uint32_t ring_dst = ring_base + 16 * lane;
  const uint8_t* ring_src = &s_ring[warp][0][0] + 12 * lane + 8;   // bytes [12*lane-8, 12*lane+20) of the segment
  uint8_t* stage = s_stage[warp];

  auto issue_row = [&](int y, int slot) {                    // y is clamped: border replicate in y
    y = min(max(y, 0), H - 1);
    if (chunk_lane)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(ring_dst + slot * RING_ROW),
                   "l"(fsrc + (size_t)((uint32_t)y * in_row)), "r"(chunk_bytes)
                   : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const uint8_t* nsrc = nullptr;                              // source of the next row to issue (clamped in y)
  auto issue_next = [&](int y, int slot) {                    // y = unclamped row of nsrc
    if (chunk_lane)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(ring_dst + slot * RING_ROW),
                   "l"(nsrc), "r"(chunk_bytes)
                   : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
    if ((unsigned)y < (unsigned)(H - 1)) nsrc += in_row;
  };
  auto fetch_row = [&](int slot, uint32_t(&r)[7]) {          // the row issued RING_D steps ago has landed
    asm volatile("cp.async.wait_group %0;" ::"n"(RING_D - 1) : "memory");
    __syncwarp();
    const volatile uint32_t* a = reinterpret_cast<const volatile uint32_t*>(ring_src + slot * RING_ROW);
#pragma unroll
    for (int k = 0; k < 7; ++k) r[k] = a[k];
  };

  u64 S[4][12];
  uint32_t raw[7];
  LaneOut o;
  {
    uint8_t* orow = out + (size_t)f * 2 * H * out_row + (size_t)xw * STAGE_ROW + (size_t)(2 * r0) * out_row;
    o.pa = orow + 16 * lane;
    o.pb = o.pa + out_row;
    o.ok1 = lane < valid_vecs;
    o.ok2 = lane < 16 && lane + 32 < valid_vecs;
  }
  const size_t adv = 2 * out_row;
  // step s handles input row q = r0 - 2 + s from ring slot s % RING_SLOTS and issues row q + RING_D
  int q = r0 - 2;
#pragma unroll
  for (int d = 0; d < RING_D; ++d) issue_row(q + d, d);
  nsrc = fsrc + (size_t)((uint32_t)min(max(q + RING_D, 0), H - 1) * in_row);
#define VEX_STEP(K, EMIT, SLOT, OK)                                   \
  {                                                                   \
  const bool ok_ = OK;                                                \
  fetch_row((SLOT) & (RING_SLOTS - 1), raw);                          \
  issue_next(q + RING_D, ((SLOT) + RING_D) & (RING_SLOTS - 1));       \
  ++q;                                                                \
  strip_step<ROUND, K, EMIT>(S, raw, left, right, stage, o, ok_, lane); \
  }
  VEX_STEP(0, false, 0, false)
  VEX_STEP(1, false, 1, false)
  VEX_STEP(2, false, 2, false)
  VEX_STEP(3, false, 3, false)
  // main loop: the step that consumes input row q emits output rows 2(q-2), 2(q-2)+1
  const int qend = min(r0 + R, H) + 2;
#pragma unroll 1
  for (; q < qend;) {
    VEX_STEP(0, true, 4, true)
    o.pa += adv;
    o.pb += adv;
    VEX_STEP(1, true, 5, q < qend)
    o.pa += adv;
    o.pb += adv;
    VEX_STEP(2, true, 6, q < qend)
    o.pa += adv;
    o.pb += adv;
    VEX_STEP(3, true, 7, q < qend)
    o.pa += adv;
    o.pb += adv;
    VEX_STEP(0, true, 0, q < qend)
    o.pa += adv;
    o.pb += adv;
    VEX_STEP(1, true, 1, q < qend)
    o.pa += adv;
    o.pb += adv;
    VEX_STEP(2, true, 2, q < qend)
    o.pa += adv;
    o.pb += adv;
    VEX_STEP(3, true, 3, q < qend)
    o.pa += adv;
    o.pb += adv;
  }
#undef VEX_STEP
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <int ROUND>
__device__ __forceinline__ uint32_t quant(float v) {
  // v = exact value * 65536.  clip to [0,255] then truncate (ROUND=0) or round half up (ROUND=1).
  float s = fmaf(v, 1.0f / 65536.0f, ROUND ? 0.5f : 0.0f);
  return min(__float2uint_rz(s), 255u);   // float->uint conversion clamps negatives to 0
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

// fp32 NCHW in [0,1] -> clip(bicubic * 255, 0, 255) truncated, u8 HWC: the tensor branch of app.py:367-376.  Plain fp32
// in ATen's order (upsample_bicubic2d: cubic_interp1d along x inside each of the 4 rows, then along y; each is
// x0*c0 + x1*c1 + x2*c2 + x3*c3 left to right, no FMA contraction).  One thread per output byte; not a throughput path.
__global__ void __launch_bounds__(256)
bicubic2x_f32in_kernel(const float* __restrict__ in, uint8_t* __restrict__ out, int n, int h, int w) {
  const size_t total = (size_t)n * h * w * 12;
  const float cw[2][4] = {{-27.f / 256.f, 225.f / 256.f, 67.f / 256.f, -9.f / 256.f},      // t = 0.25 (odd output index)
                          {-9.f / 256.f, 67.f / 256.f, 225.f / 256.f, -27.f / 256.f}};     // t = 0.75 (even)
  for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < total; j += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(j % 3);
    size_t t = j / 3;
    const int dx = (int)(t % (2 * w)); t /= 2 * w;
    const int dy = (int)(t % (2 * h));
    const int img = (int)(t / (2 * h));
    const int ix = (dx & 1) ? dx / 2 : dx / 2 - 1, iy = (dy & 1) ? dy / 2 : dy / 2 - 1;
    const float* wx = cw[(dx & 1) ? 0 : 1];
    const float* wy = cw[(dy & 1) ? 0 : 1];
    const float* plane = in + ((size_t)img * 3 + c) * h * w;
    float rows[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const float* r = plane + (size_t)min(max(iy - 1 + a, 0), h - 1) * w;
      float v = __fmul_rn(r[min(max(ix - 1, 0), w - 1)], wx[0]);
      v = __fadd_rn(v, __fmul_rn(r[min(max(ix, 0), w - 1)], wx[1]));
      v = __fadd_rn(v, __fmul_rn(r[min(max(ix + 1, 0), w - 1)], wx[2]));
      v = __fadd_rn(v, __fmul_rn(r[min(max(ix + 2, 0), w - 1)], wx[3]));
      rows[a] = v;
    }
    float v = __fmul_rn(rows[0], wy[0]);
    v = __fadd_rn(v, __fmul_rn(rows[1], wy[1]));
    v = __fadd_rn(v, __fmul_rn(rows[2], wy[2]));
    v = __fadd_rn(v, __fmul_rn(rows[3], wy[3]));
    out[j] = (uint8_t)__float2uint_rz(fminf(fmaxf(__fmul_rn(v, 255.0f), 0.f), 255.f));   // np.clip(x*255,0,255).astype(uint8)
  }
}

}  // namespace

int launch_bicubic2x_f32in(vex_ctx* ctx, const float* in, uint8_t* out, int n, int h, int w, cudaStream_t s) {
  const size_t total = (size_t)n * h * w * 12;
  const int blocks = (int)min((total + 255) / 256, (size_t)ctx->sm_count * 32);
  bicubic2x_f32in_kernel<<<blocks, 256, 0, s>>>(in, out, n, h, w);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  return VEX_OK;
}

int launch_bicubic2x(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w,
                     int round_mode, cudaStream_t s) {
  if (!in || !out || n <= 0 || h <= 0 || w <= 0 || (round_mode != 0 && round_mode != 1))
    return fail(ctx, VEX_ERR_INVALID, "vex_bicubic2x_u8: bad argument");
  const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  if (w % 16 == 0 && aligned && !ctx->opt_force_generic_bicubic) {
    const int wpr = (w + WARP_PX - 1) / WARP_PX;
    // rows per strip: tall strips amortise the 4-row window warm-up (measured: 120 rows +3 % over 40); take the
    // tallest strip that still gives every resident warp slot (12 per SM) one strip
    int R = 8;
    const long long want = (long long)ctx->sm_count * RESIDENT_WARPS;
    for (int cand : {120, 60, 40, 24, 16}) {
      if ((long long)n * ((h + cand - 1) / cand) * wpr >= want) { R = cand; break; }
    }
    if (ctx->opt_bicubic_rows > 0) R = (ctx->opt_bicubic_rows + 3) / 4 * 4;
    const int nstrips = (h + R - 1) / R;
    const long long per_frame = (long long)nstrips * wpr;
    const int max_frames = (int)std::max<long long>(1, (1LL << 30) / per_frame);
    for (int f0 = 0; f0 < n; f0 += max_frames) {
      const int nf = min(n - f0, max_frames);
      const int total_warps = (int)(per_frame * nf);
      const int blocks = (total_warps + STRIP_WARPS - 1) / STRIP_WARPS;
      const uint8_t* pin = in + (size_t)f0 * h * w * 3;
      uint8_t* pout = out + (size_t)f0 * h * w * 12;
      if (round_mode)
        bicubic2x_strip_kernel<1><<<blocks, STRIP_WARPS * 32, 0, s>>>(pin, pout, h, w, R, wpr, nstrips, total_warps);
      else
        bicubic2x_strip_kernel<0><<<blocks, STRIP_WARPS * 32, 0, s>>>(pin, pout, h, w, R, wpr, nstrips, total_warps);
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
