// The thin layers of AdvancedUpscaler on the sm_100a tensor cores (tcgen05 + TMEM), so that they run at
// HBM speed instead of CUDA-core FMA speed.  Two kernel templates:
//
//  im2col27_kernel  3-channel 3x3 neighbourhood (K = 27 -> 32) x N output channels.
//      MODE_CONV1:  uint8 HWC frame -> conv1 + bn1 + relu -> x1 bf16 NHWC             (app.py:291, 535)
//      MODE_UP1:    x5 fp32 -> ConvTranspose2d(3,64,4,2,1) + relu -> `up` bf16 NHWC    (app.py:298)
//                   as conv3x3 3 -> 4*64 followed by PixelShuffle(2) in the store addressing.
//      Builder warps gather the neighbourhood of one pixel per thread and write it as one 128 B row of a
//      K-major SWIZZLE_128B operand; one thread issues 2 tcgen05.mma (K = 2 x 16) per 128 pixels; epilogue
//      warps drain TMEM (bias, relu, bf16) to global.  TMEM holds 512/N accumulators so the three roles overlap.
//
//  gemm27_kernel    3x3 conv C -> 3 channels computed as ONE 1x1 GEMM  Y[p][(tap,co)] = sum_c X[p][c] W[c][tap][co]
//      (N = 27 -> 32) over a TMA-loaded halo tile, followed by a 9-term shifted gather-sum in shared memory.
//      A direct implicit GEMM would waste >80 % of each MMA on padding N = 3 -> 16 and re-read the tile 9 times.
//      MODE_CONV5:  x4 -> conv5 -> x5 fp32 [p][4]                                      (app.py:295)
//      MODE_UP2:    up -> ConvTranspose2d(64,3,3,1,1) -> clamp -> x255 -> truncate -> uint8 HWC (app.py:299, 549)
#pragma once

#include <cuda_bf16.h>

#include "ptx.cuh"

namespace vex {
namespace thin {

__device__ __forceinline__ uint32_t pack_f16x2_sat_thin(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ============================================================================================ im2col27
constexpr int MODE_CONV1 = 0, MODE_UP1 = 1, MODE_CONV1_F32 = 2;   // _F32: fp32 NCHW input as hi + lo bf16 (K = 64)
// MODE_FIRST64: conv_first of the Real-ESRGAN network (realesrgan_handler.py:64,87): uint8 HWC -> Conv2d(3, 64), no activation,
// fp16 NHWC out with pixel stride `out_cs`.  K = 64: the input values sit at k = 0..26 AND k = 32..58, the weight image
// holds bf16(w) at the first and bf16(w - bf16(w)) at the second position, so the weights carry 16 mantissa bits.
constexpr int MODE_FIRST64 = 3;
// MODE_CONV1_H: conv1 with fp16 operands built WITHOUT any integer->float conversion: the half 0x6400 | v is exactly
// 1024 + v for a byte v, so an operand pair is ONE byte-permute of the raw frame bytes against the constant 0x64646464.
// The GEMM then computes sum_k w_k (1024 + v_k); the weight image carries bias - 1024 * sum_k w_k (as hi + lo halves
// against two constant-1 columns), which removes the offset again - for zero-padded taps too, whose operand is 1024 + 0.
// K layout: window row ky at k = 10 ky + (3 kx + c) (9 used of 10, so no operand pair straddles two rows), k = 30, 31 bias.
// Per pixel: 9 aligned 32-bit loads + 9 funnel shifts + 15 PRMT instead of 27 byte loads + 27 I2F + 14 F2FP.
constexpr int MODE_CONV1_H = 4;
constexpr int I27_NDG = 3;             // drain groups (4 warps each) taking items k % 3: the drain is the longest role
constexpr int I27_THREADS = 32 * (9 + 4 * I27_NDG);   // warps 0-3 / 4-7 two A-builder groups, warp 8 MMA, warps 9-12 / 13-16 / 17-20 drain groups
constexpr int I27_SA = 4;              // A stages (128 rows x 128 B each); item sequence k uses stage k & 3
constexpr int I27_MMA_WARP = 8;
constexpr int I27_EPI_WARP0 = 9;
constexpr int I27_STG_PITCH = 272;     // staged row pitch: 256 B + 16 B so that 16-byte stores of 32 lanes do not conflict
constexpr int I27_STG = 32 * I27_STG_PITCH;   // per drain warp: 32 pixels x 256 B

template <int BN>
struct I27Cfg {
  static constexpr int NACC = 512 / BN;
  static constexpr int A_STAGE = 128 * 128;
  static constexpr int W_BYTES = BN * 128;
  static constexpr int W_OFF = I27_SA * A_STAGE;
  static constexpr int STG_OFF = W_OFF + W_BYTES;
  static constexpr int BAR_OFF = STG_OFF + 4 * I27_NDG * I27_STG;
  static constexpr int BIAS_OFF = BAR_OFF + 256;
  static constexpr int SMEM = BIAS_OFF + BN * 4 + 1024;
};

struct I27Params {
  const void* in;          // MODE_CONV1: uint8 [n][H][W][3];  MODE_UP1: float4 [n][H][W];  MODE_CONV1_F32: float [n][3][H][W]
  const uint8_t* wpack;    // BN rows x 128 B, SWIZZLE_128B image, k = (ky*3+kx)*3+ci for k < 27; k = 27, 28 = bias as hi + lo bf16
  __nv_bfloat16* out;      // MODE_CONV1: [n][H][W][BN];  MODE_UP1: [n][2H][2W][BN/4]
  int H, W;
  long long npix;          // n*H*W
  int in_bgr;
  int probe;               // bring-up timing probes (WRONG RESULTS): 1 = no global stores, 2 = no TMEM drain, 4 = no gather loads
  int out_cs;              // MODE_FIRST64: channel stride of the output pixels (elements); other modes: dense rows of BN
  long long* dbg;          // optional [gridDim.x][8] cycle counters (option conv_dbg): MMA thread 0 total, 1 wait acc_empty, 2 wait a_full;
                           // builder group 0: 3 wait a_empty, 4 total; drain group 0: 5 wait acc_full, 6 wait bulk-store read, 7 total
};

// 32-bit word at absolute byte address `a` (a multiple of 4) of a byte buffer [lo, hi): bytes outside the buffer read as 0
// and are never touched (the last word of a frame batch may be partial; the one before the first does not exist).
__device__ __forceinline__ uint32_t ldg_word_clipped(uintptr_t a, uintptr_t lo, uintptr_t hi) {
  if (a >= lo && a + 4 <= hi) return __ldg(reinterpret_cast<const uint32_t*>(a));
  uint32_t w = 0;
#pragma unroll
  for (int b = 0; b < 4; ++b)
    if (a + b >= lo && a + b < hi) w |= (uint32_t)__ldg(reinterpret_cast<const uint8_t*>(a + b)) << (8 * b);
  return w;
}

template <int MODE, int BN>
__global__ void __launch_bounds__(I27_THREADS, 1)
im2col27_kernel(const I27Params p) {
  using C = I27Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + C::BAR_OFF;
  const uint32_t a_full = bar0, a_empty = bar0 + 8 * I27_SA;
  const uint32_t acc_full = a_empty + 8 * I27_SA, acc_empty = acc_full + 8 * C::NACC;
  const uint32_t w_bar = acc_empty + 8 * C::NACC, tmem_slot = w_bar + 8;
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + C::BAR_OFF + (tmem_slot - bar0));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long items = (p.npix + 127) / 128;

  if (threadIdx.x == 0) {
    for (int i = 0; i < I27_SA; ++i) { ptx::mbar_init(a_full + 8 * i, 128); ptx::mbar_init(a_empty + 8 * i, 1); }
    for (int i = 0; i < C::NACC; ++i) { ptx::mbar_init(acc_full + 8 * i, 1); ptx::mbar_init(acc_empty + 8 * i, 4); }
    ptx::mbar_init(w_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == I27_MMA_WARP) {
    ptx::tmem_alloc(tmem_slot, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;

  if (warp < 8) {
    // ================================ A builders: one pixel (= one 128 B operand row) per thread.
    // The gather is latency-bound (9-27 dependent-free loads, then a barrier), so two groups of 128 threads
    // build alternate items: group gi owns item sequence numbers k = gi, gi+2, ... and stages gi, gi+2.
    const int gi = warp >> 2;
    const int r = threadIdx.x & 127;                 // row within the stage
    uint32_t sa = gi, pa = 0;
    const bool dbg = p.dbg != nullptr && threadIdx.x == 0;
    long long tbw = 0;
    const long long tb0 = dbg ? clock64() : 0;
    // ---- MODE_CONV1_H: the 3 x 9 window bytes of a pixel as 9 aligned 32-bit words (3 per row), fetched one item ahead
    const uintptr_t lo = reinterpret_cast<uintptr_t>(p.in), hi = lo + (uintptr_t)p.npix * 3;
    auto fetch_h = [&](long long pix, uint32_t (&raw)[9]) {
#pragma unroll
      for (int j = 0; j < 9; ++j) raw[j] = 0u;
      if (pix >= p.npix) return;
      const int y = (int)((pix / p.W) % p.H);
      const uintptr_t a0 = lo + (uintptr_t)((pix - 1) * 3);               // byte address of pixel (y, x-1), channel 0
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const int yy = y + ky - 1;
        if (yy < 0 || yy >= p.H) continue;                                // rows of padding stay zero
        const uintptr_t wa = (a0 + (uintptr_t)((long long)(ky - 1) * p.W * 3)) & ~(uintptr_t)3;
#pragma unroll
        for (int j = 0; j < 3; ++j) raw[3 * ky + j] = ldg_word_clipped(wa + 4 * j, lo, hi);
      }
    };
    auto pack_h = [&](long long pix, const uint32_t (&raw)[9], uint32_t (&w)[16]) {
      if (pix >= p.npix) return;
      const int x = (int)(pix % p.W);
      const uintptr_t a0 = lo + (uintptr_t)((pix - 1) * 3);
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const uint32_t sh = (uint32_t)((a0 + (uintptr_t)((long long)(ky - 1) * p.W * 3)) & 3) * 8;
        uint32_t A0 = __funnelshift_r(raw[3 * ky], raw[3 * ky + 1], sh);   // the row's 9 window bytes, aligned
        uint32_t A1 = __funnelshift_r(raw[3 * ky + 1], raw[3 * ky + 2], sh);
        uint32_t A2 = raw[3 * ky + 2] >> sh;
        if (x == 0) A0 &= 0xFF000000u;                                     // left neighbour is padding (bytes 0-2)
        if (x == p.W - 1) { A1 &= 0x0000FFFFu; A2 = 0u; }                  // right neighbour is padding (bytes 6-8)
        constexpr uint32_t C64 = 0x64646464u;
        w[5 * ky + 0] = __byte_perm(A0, C64, 0x4140);
        w[5 * ky + 1] = __byte_perm(A0, C64, 0x4342);
        w[5 * ky + 2] = __byte_perm(A1, C64, 0x4140);
        w[5 * ky + 3] = __byte_perm(A1, C64, 0x4342);
        w[5 * ky + 4] = __byte_perm(A2, C64, 0x4440);                      // k = 10 ky + 9 is unused (zero weight)
      }
      w[15] = 0x3C003C00u;                                                 // k = 30, 31: 1.0 against the bias halves
    };
    uint32_t raw[9];
    if (MODE == MODE_CONV1_H) {
      const long long it0 = blockIdx.x + (long long)gi * gridDim.x;
      if (it0 < items) fetch_h(it0 * 128 + r, raw);
    }
    for (long long it = blockIdx.x + (long long)gi * gridDim.x; it < items; it += 2LL * gridDim.x) {
      const long long pix = it * 128 + r;
      uint32_t w[16], wlo[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) w[j] = wlo[j] = 0u;
      if (MODE == MODE_CONV1_H) {
        pack_h(pix, raw, w);
        const long long itn = it + 2LL * gridDim.x;          // the raw words are consumed: the next item's loads fly during the hand-off
        if (itn < items) fetch_h(itn * 128 + r, raw);
      } else if (pix < p.npix && !(p.probe & 4)) {
        const int x = (int)(pix % p.W);
        const long long t = pix / p.W;
        const int y = (int)(t % p.H);
        const long long img = t / p.H;
        float v[27];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const int yy = y + ky - 1;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const int xx = x + kx - 1;
            const bool ok = yy >= 0 && yy < p.H && xx >= 0 && xx < p.W;
            const long long q = (img * p.H + yy) * p.W + xx;
            if (MODE == MODE_CONV1_F32) {
              const size_t plane = (size_t)p.H * p.W;
              const float* s = reinterpret_cast<const float*>(p.in) + (size_t)img * 3 * plane + (size_t)yy * p.W + xx;
#pragma unroll
              for (int c = 0; c < 3; ++c) v[(ky * 3 + kx) * 3 + c] = ok ? __ldg(s + c * plane) : 0.f;
            } else if (MODE == MODE_CONV1 || MODE == MODE_FIRST64) {
              const uint8_t* s = reinterpret_cast<const uint8_t*>(p.in) + q * 3;
#pragma unroll
              for (int c = 0; c < 3; ++c)
                v[(ky * 3 + kx) * 3 + c] = ok ? (float)__ldg(s + (p.in_bgr ? 2 - c : c)) : 0.f;
            } else {
              float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
              if (ok) f = __ldg(reinterpret_cast<const float4*>(p.in) + q);
              v[(ky * 3 + kx) * 3 + 0] = f.x; v[(ky * 3 + kx) * 3 + 1] = f.y; v[(ky * 3 + kx) * 3 + 2] = f.z;
            }
          }
        }
#pragma unroll
        for (int j = 0; j < 13; ++j) w[j] = ptx::pack_bf16x2(v[2 * j], v[2 * j + 1]);
        w[13] = ptx::pack_bf16x2(v[26], 1.0f);       // k = 27, 28: constant 1 against the bias rows (hi + lo bf16) of W, so the
        w[14] = ptx::pack_bf16x2(1.0f, 0.f);         // drain needs no per-element bias load and add
        if (MODE == MODE_FIRST64) {                  // k = 32..58: the same (exact) input values against the lo halves of W
#pragma unroll
          for (int j = 0; j < 13; ++j) wlo[j] = w[j];
          wlo[13] = ptx::pack_bf16x2(v[26], 0.f);
        }
        if (MODE == MODE_CONV1_F32) {                // k = 32..58: what the bf16 rounding of k = 0..26 dropped
#pragma unroll
          for (int j = 0; j < 27; ++j) {
            const uint32_t hw = (j & 1) ? (w[j >> 1] & 0xFFFF0000u) : (w[j >> 1] << 16);
            v[j] -= __uint_as_float(hw);
          }
#pragma unroll
          for (int j = 0; j < 13; ++j) wlo[j] = ptx::pack_bf16x2(v[2 * j], v[2 * j + 1]);
          wlo[13] = ptx::pack_bf16x2(v[26], 0.f);
        }
      }
      const long long tq0 = dbg ? clock64() : 0;
      ptx::mbar_wait(a_empty + 8 * sa, pa ^ 1);
      if (dbg) tbw += clock64() - tq0;
      uint8_t* row = sm + sa * C::A_STAGE + r * 128;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(row + ((j ^ (r & 7)) << 4)) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
      if (MODE == MODE_CONV1_F32 || MODE == MODE_FIRST64) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(row + (((4 + j) ^ (r & 7)) << 4)) = make_uint4(wlo[4 * j], wlo[4 * j + 1], wlo[4 * j + 2], wlo[4 * j + 3]);
      }
      ptx::fence_proxy_async_smem();                 // generic-proxy stores -> visible to the tensor core
      ptx::mbar_arrive(a_full + 8 * sa);
      sa += 2;
      if (sa >= I27_SA) { sa -= I27_SA; pa ^= 1; }
    }
    if (dbg) { p.dbg[blockIdx.x * 8 + 3] = tbw; p.dbg[blockIdx.x * 8 + 4] = clock64() - tb0; }
  } else if (warp == I27_MMA_WARP) {
    // ================================ weights once, then 2 MMAs per item
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(w_bar, C::W_BYTES);
      ptx::bulk_load(sbase + C::W_OFF, p.wpack, C::W_BYTES, w_bar);
      ptx::mbar_wait(w_bar, 0);
      // kind::f16 instruction descriptor: bits 7-9 / 10-12 = A / B format (0 fp16, 1 bf16)
      constexpr uint32_t idesc = MODE == MODE_CONV1_H ? (ptx::idesc_bf16_f32(128, BN) & ~((1u << 7) | (1u << 10)))
                                                      : ptx::idesc_bf16_f32(128, BN);
      uint32_t sa = 0, pa = 0, b = 0, pb = 0;
      const bool dbg = p.dbg != nullptr;
      long long t1 = 0, t2 = 0, tq = 0;
      const long long tm0 = dbg ? clock64() : 0;
      for (long long it = blockIdx.x; it < items; it += gridDim.x) {
        if (dbg) tq = clock64();
        ptx::mbar_wait(acc_empty + 8 * b, pb ^ 1);
        if (dbg) { const long long n = clock64(); t1 += n - tq; tq = n; }
        ptx::mbar_wait(a_full + 8 * sa, pa);
        if (dbg) t2 += clock64() - tq;
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < ((MODE == MODE_CONV1_F32 || MODE == MODE_FIRST64) ? 4 : 2); ++k) {
          const uint64_t ad = ptx::smem_desc_sw128(sbase + sa * C::A_STAGE + k * 32, 0);
          const uint64_t bd = ptx::smem_desc_sw128(sbase + C::W_OFF + k * 32, 0);
          ptx::mma_bf16_ss(tmem + b * BN, ad, bd, idesc, k);
        }
        ptx::tc_commit(a_empty + 8 * sa);
        ptx::tc_commit(acc_full + 8 * b);
        if (++sa == I27_SA) { sa = 0; pa ^= 1; }
        if (++b == C::NACC) { b = 0; pb ^= 1; }
      }
      if (dbg) { long long* d = p.dbg + blockIdx.x * 8; d[0] = clock64() - tm0; d[1] = t1; d[2] = t2; }
    }
  } else {
    // ================================ drain: TMEM -> bias, relu, bf16 -> shared stage -> bulk (TMA) stores.
    // Two groups of 4 warps take alternate items.  A lane owns one pixel (TMEM lane); per 128 output columns it
    // stages ITS OWN 256-byte row (pitch 272 B: conflict-free 16-byte stores) and sends it to global with one
    // cp.async.bulk of 256 contiguous bytes:
    //   CONV1: channels [128*hf, 128*hf+128) of the pixel;
    //   UP1:   columns [128*py, 128*py+128) = hi-res pixels (2y+py, 2x) and (2y+py, 2x+1), 64 channels each (PixelShuffle).
    // No lane touches another lane's row, so there is no warp synchronisation and no address shuffling in the loop.
    const int ew = warp - I27_EPI_WARP0;
    const int ge = ew >> 2, q = warp & 3;
    uint8_t* stg = sm + C::STG_OFF + ew * I27_STG + lane * I27_STG_PITCH;
    const uint32_t stg_s = ptx::smem_u32(stg);
    constexpr int HALVES = BN >= 128 ? BN / 128 : 1;
    constexpr int GPH = BN >= 128 ? 4 : BN / 32;   // 32-column groups per staged row (row = GPH * 64 bytes)
    long long k = ge;
    const bool dbg = p.dbg != nullptr && ew == 0 && lane == 0;
    long long t5 = 0, t6 = 0;
    const long long td0 = dbg ? clock64() : 0;
    // A parity wait is only safe for a waiter that is at most one phase behind its barrier, so the groups that take items
    // in turn must not outnumber the accumulators: with 2 accumulators (BN = 256) a third group's first item would be the
    // SECOND use of accumulator 0, its wait for parity 1 would pass on the untouched barrier and it would drain garbage.
    constexpr int NDG = C::NACC < I27_NDG ? C::NACC : I27_NDG;
    for (long long it = ge < NDG ? blockIdx.x + (long long)ge * gridDim.x : items; it < items;
         it += (long long)NDG * gridDim.x, k += NDG) {
      const uint32_t b = (uint32_t)(k % C::NACC), pb = (uint32_t)((k / C::NACC) & 1);
      const long long pix = it * 128 + q * 32 + lane;
      const bool valid = pix < p.npix;
      uint8_t* obase = nullptr;                      // first of this pixel's HALVES 256-byte destinations
      size_t half_stride = 256;                      // CONV1: next 128 channels of the same pixel
      if (valid) {
        if (MODE == MODE_FIRST64) {
          obase = reinterpret_cast<uint8_t*>(p.out + pix * p.out_cs);
        } else if (MODE != MODE_UP1) {
          obase = reinterpret_cast<uint8_t*>(p.out + pix * BN);
        } else {
          const int x = (int)(pix % p.W);
          const long long t = pix / p.W;
          const int y = (int)(t % p.H);
          const long long img = t / p.H;
          obase = reinterpret_cast<uint8_t*>(p.out + ((img * 2 * p.H + 2 * y) * (2LL * p.W) + 2 * x) * (BN / 4));
          half_stride = (size_t)(2LL * p.W) * (BN / 4) * 2;     // hi-res row 2y+1
        }
      }
      const long long tq5 = dbg ? clock64() : 0;
      ptx::mbar_wait(acc_full + 8 * b, pb);
      if (dbg) t5 += clock64() - tq5;
      ptx::tc_fence_after();
#pragma unroll 1
      for (int hf = 0; hf < HALVES; ++hf) {
        const long long tq6 = dbg ? clock64() : 0;
        ptx::bulk_wait_read0();                      // this lane's previous bulk store has finished reading its row
        if (dbg) t6 += clock64() - tq6;
#pragma unroll 1
        for (int g = (p.probe & 2) ? GPH : 0; g < GPH; ++g) {
          uint32_t v[32];
          const int col = hf * 128 + g * 32;
          ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + b * BN + col, v);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[j4 * 8 + e]);   // bias already inside (k = 27, 28)
            if (MODE == MODE_FIRST64)                // linear, fp16 (the RRDB trunk runs in fp16)
              *reinterpret_cast<uint4*>(stg + (g * 4 + j4) * 16) =
                  make_uint4(pack_f16x2_sat_thin(f[0], f[1]), pack_f16x2_sat_thin(f[2], f[3]),
                             pack_f16x2_sat_thin(f[4], f[5]), pack_f16x2_sat_thin(f[6], f[7]));
            else
            *reinterpret_cast<uint4*>(stg + (g * 4 + j4) * 16) =
                make_uint4(ptx::pack_relu_bf16x2(f[0], f[1]), ptx::pack_relu_bf16x2(f[2], f[3]),
                           ptx::pack_relu_bf16x2(f[4], f[5]), ptx::pack_relu_bf16x2(f[6], f[7]));
          }
        }
        if (hf == HALVES - 1) {                      // accumulator fully read: hand it back before the stores
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(acc_empty + 8 * b);
        }
        ptx::fence_proxy_async_smem();               // the row was written through the generic proxy
        if (p.probe & 8) {                           // probe: half as many ops, 512 B each (garbage payload)
          if (valid && !(lane & 1)) ptx::bulk_store(obase + hf * half_stride, stg_s, 512);
        } else if (valid && !(p.probe & 1)) ptx::bulk_store(obase + hf * half_stride, stg_s, GPH * 64);
        ptx::bulk_commit();
      }
    }
    ptx::bulk_wait_all0();                           // global writes complete before the CTA exits
    if (dbg) { long long* d = p.dbg + blockIdx.x * 8; d[5] = t5; d[6] = t6; d[7] = clock64() - td0; }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == I27_MMA_WARP) {
    __syncwarp();
    ptx::tmem_dealloc(tmem, 512);
  }
}

// ============================================================================================== gemm27
constexpr int MODE_CONV5 = 0, MODE_UP2 = 1;
// MODE_LAST_F16: conv_last of the Real-ESRGAN network (realesrgan_handler.py:78,99 + the clamp / x255 / truncating cast of
// :318-319): fp16 NHWC in, fp16 weight image, otherwise MODE_UP2's epilogue.
constexpr int MODE_LAST_F16 = 2;
constexpr int G27_THREADS = 320;       // warp 0 TMA, warp 1 MMA, warps 2-5 / 6-9 two epilogue groups (even / odd items)
constexpr int G27_ROWS = 6;            // output rows per item; halo tile = 8 rows x 32 px = 256 GEMM rows
constexpr int G27_TW = 30;
constexpr int G27_SA = 4;
constexpr int G27_A_STAGE = 8 * 32 * 128;            // 32 KB: one 64-channel chunk of the halo tile
constexpr int G27_YS_PITCH = 256;                    // Ys[k][m], k < 27, m = halo-flat pixel
constexpr int G27_OB_PITCH = 104;                    // staged uint8 output row (<= 3 + 90 bytes), word aligned

struct G27Cfg {
  static constexpr int W_OFF = G27_SA * G27_A_STAGE;               // up to 8 chunks x 4 KB of weights
  static constexpr int YS_OFF = W_OFF + 8 * 4096;
  static constexpr int YS_BYTES = 27 * G27_YS_PITCH * 4;            // per epilogue group
  static constexpr int OB_OFF = YS_OFF + 2 * YS_BYTES;
  static constexpr int OB_BYTES = G27_ROWS * G27_OB_PITCH + 16;     // per epilogue group
  static constexpr int BAR_OFF = OB_OFF + 2 * OB_BYTES;
  static constexpr int SMEM = BAR_OFF + 256 + 1024;
};

struct G27Params {
  const uint8_t* wpack;    // [C/64] slices of 32 rows x 128 B (SWIZZLE_128B image); row n = tap*3+co
  const float* bias;       // [3]
  float4* out_f4;          // MODE_CONV5: [n][H][W]
  uint8_t* out_u8;         // MODE_UP2:   [n][H][W][3]
  float* out_nchw;         // MODE_UP2, optional: [n][3][H][W] clamp(...) before quantisation
  int H, W, NC;            // NC = C / 64
  int tiles_x, tiles_y, n_img;
  int out_bgr;
};

template <int MODE>
__global__ void __launch_bounds__(G27_THREADS, 1)
gemm27_kernel(const __grid_constant__ CUtensorMap amap, const G27Params p) {
  using C = G27Cfg;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + C::BAR_OFF;
  const uint32_t a_full = bar0, a_empty = bar0 + 8 * G27_SA;
  const uint32_t acc_full = a_empty + 8 * G27_SA, acc_empty = acc_full + 16;
  const uint32_t w_bar = acc_empty + 16, tmem_slot = w_bar + 8;
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + C::BAR_OFF + (tmem_slot - bar0));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = p.tiles_x * p.tiles_y * p.n_img;
  // The epilogue (TMEM drain, 9-tap gather, quantise, row stores) is a long dependent chain with three barriers;
  // two groups of 4 warps work on alternate items, each with its own TMEM buffer, Ys planes and output stage.
  const int grp = warp >= 6 ? 1 : 0;
  float* Ys = reinterpret_cast<float*>(sm + C::YS_OFF + grp * C::YS_BYTES);
  uint8_t* ob = sm + C::OB_OFF + grp * C::OB_BYTES;

  if (threadIdx.x == 0) {
    for (int i = 0; i < G27_SA; ++i) { ptx::mbar_init(a_full + 8 * i, 1); ptx::mbar_init(a_empty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(acc_full + 8 * i, 1); ptx::mbar_init(acc_empty + 8 * i, 4); }
    ptx::mbar_init(w_bar, 1);
    ptx::fence_mbar_init();
    ptx::prefetch_tensormap(&amap);
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 128);                 // 2 buffers x (2 groups x 32 columns)
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;

  if (warp == 0) {
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(w_bar, p.NC * 4096);
      ptx::bulk_load(sbase + C::W_OFF, p.wpack, p.NC * 4096, w_bar);
      uint32_t sa = 0, pa = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x) {
        const int txi = it % p.tiles_x, r = it / p.tiles_x;
        const int tyi = r % p.tiles_y, n = r / p.tiles_y;
        for (int c = 0; c < p.NC; ++c) {
          ptx::mbar_wait(a_empty + 8 * sa, pa ^ 1);
          ptx::mbar_expect_tx(a_full + 8 * sa, G27_A_STAGE);
          ptx::tma_load_4d(sbase + sa * G27_A_STAGE, &amap, a_full + 8 * sa, c * 64, txi * G27_TW - 1,
                           tyi * G27_ROWS - 1, n);
          if (++sa == G27_SA) { sa = 0; pa ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (ptx::elect_one()) {
      ptx::mbar_wait(w_bar, 0);
      // kind::f16 instruction descriptor: bits 7-9 / 10-12 = A / B format (0 fp16, 1 bf16)
      constexpr uint32_t idesc = MODE == MODE_LAST_F16 ? (ptx::idesc_bf16_f32(128, 32) & ~((1u << 7) | (1u << 10)))
                                                       : ptx::idesc_bf16_f32(128, 32);
      uint32_t sa = 0, pa = 0, b = 0, pb = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x) {
        ptx::mbar_wait(acc_empty + 8 * b, pb ^ 1);
        ptx::tc_fence_after();
        for (int c = 0; c < p.NC; ++c) {
          ptx::mbar_wait(a_full + 8 * sa, pa);
          ptx::tc_fence_after();
#pragma unroll
          for (int g = 0; g < 2; ++g)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t ad = ptx::smem_desc_sw128(sbase + sa * G27_A_STAGE + g * 16384 + k * 32, 0);
              const uint64_t bd = ptx::smem_desc_sw128(sbase + C::W_OFF + c * 4096 + k * 32, 0);
              ptx::mma_bf16_ss(tmem + b * 64 + g * 32, ad, bd, idesc, (c | k) != 0);
            }
          ptx::tc_commit(a_empty + 8 * sa);
          if (++sa == G27_SA) { sa = 0; pa ^= 1; }
        }
        ptx::tc_commit(acc_full + 8 * b);
        if (++b == 2) { b = 0; pb ^= 1; }
      }
    }
  } else {
    const int q = warp & 3;
    const int et = ((warp - 2) & 3) * 32 + lane;     // 0..127 among the threads of this epilogue group
    const float b0 = p.bias[0], b1 = p.bias[1], b2 = p.bias[2];
    const uint32_t b = grp;                          // item sequence parity = TMEM buffer = group
    uint32_t pb = 0;
    for (int it = blockIdx.x + grp * gridDim.x; it < total; it += 2 * gridDim.x) {
      const int txi = it % p.tiles_x, r = it / p.tiles_x;
      const int tyi = r % p.tiles_y, n = r / p.tiles_y;
      const int x0 = txi * G27_TW, y0 = tyi * G27_ROWS;
      ptx::mbar_wait(acc_full + 8 * b, pb);
      ptx::tc_fence_after();
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + b * 64 + g * 32, v);
        ptx::tmem_ld_wait();
        const int m = g * 128 + q * 32 + lane;
#pragma unroll
        for (int k = 0; k < 27; ++k) Ys[k * G27_YS_PITCH + m] = __uint_as_float(v[k]);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc_empty + 8 * b);      // TMEM buffer free: next-but-one item may start
      pb ^= 1;
      named_bar_sync(1 + grp, 128);
      // 9-tap gather-sum: out(ty,tx) = bias + sum_t Y[(ty+t/3)*32 + tx+t%3][t*3+co]
      const int nvx = min(G27_TW, p.W - x0);                    // valid pixels per row of this tile
#pragma unroll 1
      // one warp = one tile row (lanes 30, 31 idle): a warp's 32 loads then fall into one row of Ys, i.e. 32 banks
      for (int idx = et; idx < G27_ROWS * 32; idx += 128) {
        const int ty = idx >> 5, tx = idx & 31;
        if (tx >= G27_TW) continue;
        const int y = y0 + ty, x = x0 + tx;
        float s0 = b0, s1 = b1, s2 = b2;
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const int m = (ty + t / 3) * 32 + tx + t % 3;
          s0 += Ys[(t * 3 + 0) * G27_YS_PITCH + m];
          s1 += Ys[(t * 3 + 1) * G27_YS_PITCH + m];
          s2 += Ys[(t * 3 + 2) * G27_YS_PITCH + m];
        }
        if (y < p.H && x < p.W) {
          const size_t pix = ((size_t)n * p.H + y) * p.W + x;
          if (MODE == MODE_CONV5) {
            p.out_f4[pix] = make_float4(s0, s1, s2, 0.f);
          } else {
            const float c0 = fminf(fmaxf(s0, 0.f), 1.f), c1 = fminf(fmaxf(s1, 0.f), 1.f), c2 = fminf(fmaxf(s2, 0.f), 1.f);
            if (p.out_nchw) {
              const size_t plane = (size_t)p.H * p.W;
              float* o = p.out_nchw + (size_t)n * 3 * plane + (size_t)y * p.W + x;
              o[0] = c0; o[plane] = c1; o[2 * plane] = c2;
            }
            if (p.out_u8) {
              // np.clip(x*255, 0, 255).astype(np.uint8): fp32 product, truncation (app.py:549)
              const uint8_t q0 = (uint8_t)__float2uint_rz(c0 * 255.0f), q1 = (uint8_t)__float2uint_rz(c1 * 255.0f),
                            q2 = (uint8_t)__float2uint_rz(c2 * 255.0f);
              // staged with the same (address & 3) phase as the global row so that words line up
              const size_t rowb = (((size_t)n * p.H + y) * p.W + x0) * 3;
              const int ph = (int)(reinterpret_cast<uintptr_t>(p.out_u8 + rowb) & 3);
              uint8_t* s = ob + ty * G27_OB_PITCH + ph + tx * 3;
              s[0] = p.out_bgr ? q2 : q0; s[1] = q1; s[2] = p.out_bgr ? q0 : q2;
            }
          }
        }
      }
      if (MODE != MODE_CONV5 && p.out_u8) {
        named_bar_sync(1 + grp, 128);
        // cooperative row stores: aligned 32-bit words in the middle, bytes at the ragged ends
        for (int wi = et; wi < G27_ROWS * 24; wi += 128) {
          const int ty = wi / 24, j = wi - ty * 24;
          const int y = y0 + ty;
          if (y >= p.H) continue;
          const size_t rowb = (((size_t)n * p.H + y) * p.W + x0) * 3;
          const int ph = (int)(reinterpret_cast<uintptr_t>(p.out_u8 + rowb) & 3), nbytes = nvx * 3;
          const int lo = j * 4, hi = lo + 4;                   // staged byte range of this word
          if (lo >= ph + nbytes || hi <= ph) continue;
          uint8_t* g = p.out_u8 + (rowb - ph) + lo;
          const uint8_t* s = ob + ty * G27_OB_PITCH + lo;
          if (lo >= ph && hi <= ph + nbytes) {
            *reinterpret_cast<uint32_t*>(g) = *reinterpret_cast<const uint32_t*>(s);
          } else {
            for (int e = 0; e < 4; ++e)
              if (lo + e >= ph && lo + e < ph + nbytes) g[e] = s[e];
          }
        }
      }
      named_bar_sync(1 + grp, 128);                             // Ys / ob reusable
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    ptx::tmem_dealloc(tmem, 128);
  }
}

}  // namespace thin
}  // namespace vex
