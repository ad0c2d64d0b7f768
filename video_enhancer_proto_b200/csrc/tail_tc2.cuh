// Fused tail of AdvancedUpscaler, second generation: x5 -> upconv1 + relu -> upconv2 -> clamp -> uint8 (app.py:298-299, 549)
// with the `up` tensor kept in TENSOR MEMORY between the two GEMMs.
//
// The first generation (tail_tc.cuh) wrote relu(upconv1) as the bf16 A operand of the second GEMM into shared memory, in
// PixelShuffle order (a [256 hi-res px x 64 ch] tile).  Its per-role cycle counters (option conv_dbg) showed, once the x5
// windows came by TMA, the mid-drain warps busy 1425 of 1960 cycles per item with only 94 of them in tcgen05.ld: the
// kernel is bound by the 128 B/clk shared-memory pipe (A2 written: 49 KB and read by the MMA: 32 KB per item, next to
// the Ys planes, the operands of the first GEMM and the x5 windows: ~1550 wavefront-cycles per item).
//
// Here the PixelShuffle is not materialised at all.  The hi-res image is treated as FOUR phase planes (py, px), each
// indexed by the low-res pixel: A2_ph[r][ch] = relu(D1[r][ph*64 + ch]) lives in the SAME TMEM lane r as the accumulator row
// it comes from, so the mid-drain is tcgen05.ld -> relu/bf16 pack -> tcgen05.st, with no shared memory, no swizzle and no
// cross-lane movement, and the second GEMM takes its A operand from TMEM (tcgen05.mma [d], [a], b-desc; layout measured
// with tools/ubench/mma_ts.cu: lane = row, column j = elements 2j | 2j+1, a K = 16 step = 8 columns):
//   Y_ph[r][(tap,co)] = sum_ch A2_ph[r][ch] * W2[ch][tap][co]           4 phases x (M = 128, N = 32, K = 64)
// The 9-tap gather of the epilogue finds hi-res neighbour (hr, hc) of the halo tile in plane ((hr+1)&1, (hc+1)&1) at low-res
// pixel ((hr+1)>>1, (hc+1)>>1); the planes are 112 words apart so that a warp's 30 loads fall into 32 different banks.
//
// TMEM columns: [0,256) first accumulator (two halves = phase pairs), [256,384) A2 (4 x 32: 64 bf16 each), [384,512) Y (4 x 32).
// Results are bit-identical to tail_tc.cuh (same products, same K order per accumulator, same tap order in the gather;
// tests/test_cnn_gpu.py::test_fused_tail_generations_bit_identical).
//
// Pipelining.  The 512 TMEM columns are all taken (256 + 128 + 128), but every buffer exists once PER HALF (phase pair):
// unit (item k, half h) flows GEMM #1 -> acc1[h] -> mid-drain group h -> A2[h] -> GEMM #2 -> Y[h] -> epilogue, and the two
// halves are independent pipelines.  The first version issued all four GEMMs of an item from ONE thread in program order
// (GEMM #2 of item k-1 for both halves, then GEMM #1 of item k): blocked on one half's barrier it could not serve the
// other, everything ran in series and an item took 2287 cycles against 1961 for tail_tc.cuh (option conv_dbg: the MMA
// thread waited 999 cycles per item for the A2 planes).  Now each half has its OWN issuing thread (warp 7 / warp 3), the
// mid-drain hands the accumulator back as soon as it has been read (before the last store), and only 3 builder warps remain
// (85 live patch rows): 2007-2022 cycles per item.
//
// STATUS: a measured variant (option fuse_tail = 3), not the default - it ties with tail_tc.cuh (1960 cycles per item) and
// does not beat it.  The counters say why: the mid-drain warps are still "busy" ~1360 cycles per item although their
// tcgen05.ld take 108 and their tcgen05.st 96 of them and they touch no shared memory at all.  What is left is their
// instruction stream itself (64 conversions + 64 selects + addressing per thread and item) issued at ~0.25 instructions
// per clock, because 24 warps share four schedulers that are 58 % busy (ncu: issue active 58 %, ALU 48 %).  The tail is
// bound by warp-instruction issue across its five roles, not by TMEM, not by shared memory alone; the next lever is fewer
// instructions per output pixel (a 16 x 16 halo tile: 196 instead of 180 outputs for the same work; the 3-tap row sums of
// the gather by warp shuffles), not another data path.
#pragma once

#include "tail_tc.cuh"

namespace vex {
namespace thin {

constexpr int UT2_PLANE = 112;                         // words between phase planes of Ys: >= 96 live lanes, = 16 mod 32
constexpr int UT2_YS_PITCH = 4 * UT2_PLANE;            // Ys[k][plane * 112 + r], k < 27

struct UT2Cfg {
  static constexpr int A1_STAGE = 128 * 128;
  static constexpr int A1_OFF = 0;                                  // 2 stages
  static constexpr int W1_OFF = A1_OFF + 2 * A1_STAGE;              // 256 rows x 128 B
  static constexpr int W2_OFF = W1_OFF + 256 * 128;                 // 32 rows x 128 B
  static constexpr int YS_OFF = W2_OFF + 4096;
  static constexpr int YS_BYTES = 27 * UT2_YS_PITCH * 4;            // 48,384 per epilogue group
  static constexpr int OB_OFF = YS_OFF + UT_EPI_GROUPS * YS_BYTES;
  static constexpr int OB_BYTES = G27_ROWS * G27_OB_PITCH + 16;
  static constexpr int X_OFF = (OB_OFF + UT_EPI_GROUPS * OB_BYTES + 127) / 128 * 128;
  static constexpr int BAR_OFF = X_OFF + UT_XS * UT_X_STAGE;
  static constexpr int SMEM = BAR_OFF + 256 + 1024;
};

// D[tmem] (+)= A[tmem] * B[smem]^T: the A operand (M = 128 rows x 16 bf16) is read from tensor memory
__device__ __forceinline__ void mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__global__ void __launch_bounds__(UT_THREADS, 1)
uptail2_kernel(const __grid_constant__ CUtensorMap xmap, const UTParams p) {
  using C = UT2Cfg;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + C::BAR_OFF;
  const uint32_t a1_full = bar0, a1_empty = bar0 + 16;
  const uint32_t acc1_full = bar0 + 32, acc1_empty = bar0 + 48;      // per half of the first accumulator (phases 0-1 / 2-3)
  const uint32_t a2_full = bar0 + 64, a2_empty = bar0 + 80;          // per half of the A2 planes in TMEM
  // Y is single-buffered, but each epilogue group has ITS OWN "full" barrier (completing once per item of that group): a
  // parity wait is only safe for a waiter at most one phase behind, and three groups taking turns on one barrier are not.
  const uint32_t y_full = bar0 + 96, y_empty = bar0 + 96 + 16 * UT_EPI_GROUPS;     // y_full[group][half], y_empty[half]
  const uint32_t w_bar = bar0 + 160, tmem_slot = bar0 + 168;
  const uint32_t x_full = bar0 + 176, x_empty = bar0 + 176 + 8 * UT_XS;
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + C::BAR_OFF + 168);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = p.tiles_x * p.tiles_y * p.n_img;
  const int H2 = 2 * p.H, W2 = 2 * p.W;
  constexpr uint32_t TM_A2 = 256, TM_Y = 384;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(a1_full + 8 * i, 96); ptx::mbar_init(a1_empty + 8 * i, 2);      // 3 builder warps; both issuing threads
      ptx::mbar_init(a2_full + 8 * i, 3); ptx::mbar_init(a2_empty + 8 * i, 1);
      ptx::mbar_init(acc1_full + 8 * i, 1); ptx::mbar_init(acc1_empty + 8 * i, 3);
      ptx::mbar_init(y_empty + 8 * i, 3);
    }
    for (int i = 0; i < 2 * UT_EPI_GROUPS; ++i) ptx::mbar_init(y_full + 8 * i, 1);
    ptx::mbar_init(w_bar, 1);
    for (int i = 0; i < UT_XS; ++i) { ptx::mbar_init(x_full + 8 * i, 1); ptx::mbar_init(x_empty + 8 * i, 96); }
    ptx::fence_mbar_init();
    ptx::prefetch_tensormap(&xmap);
  }
  if (warp == UT_MMA_WARP) {
    ptx::tmem_alloc(tmem_slot, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;

  if (warp < 3) {
    // ================================ builders: low-res pixel r of the 5 x 17 patch -> one operand row (as tail_tc.cuh);
    // rows 96-127 of the operand are never written and the accumulator lanes they feed are never read
    const int r = threadIdx.x;
    const int lr = r / UT_LC, lc = r - lr * UT_LC;
    uint32_t k = 0;
    const bool dbg = p.dbg != nullptr && threadIdx.x == 0;
    long long tw = 0, tx = 0, tb0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
      const int txi = it % p.tiles_x, t = it / p.tiles_x;
      const int tyi = t % p.tiles_y;
      const int y = 3 * tyi - 1 + lr, x = 15 * txi - 1 + lc;
      const bool live = r < UT_ROWS1 && y >= 0 && y < p.H && x >= 0 && x < p.W;
      const uint32_t xs = k % UT_XS, xp = (k / UT_XS) & 1;
      uint32_t w[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) w[j] = 0u;
      const long long tx0 = dbg ? clock64() : 0;
      ptx::mbar_wait(x_full + 8 * xs, xp);
      if (dbg) tx += clock64() - tx0;
      if (live) {
        const float4* win = reinterpret_cast<const float4*>(sm + C::X_OFF + xs * UT_X_STAGE) + lr * UT_XW + lc;
        float v[27];
#pragma unroll
        for (int q9 = 0; q9 < 9; ++q9) {
          const float4 f = win[(q9 / 3) * UT_XW + q9 % 3];
          v[q9 * 3 + 0] = f.x; v[q9 * 3 + 1] = f.y; v[q9 * 3 + 2] = f.z;
        }
#pragma unroll
        for (int j = 0; j < 13; ++j) w[j] = ptx::pack_bf16x2(v[2 * j], v[2 * j + 1]);
        w[13] = ptx::pack_bf16x2(v[26], 1.0f);       // k = 27, 28: constant 1 against the bias rows of W1
        w[14] = ptx::pack_bf16x2(1.0f, 0.f);
      }
      ptx::mbar_arrive(x_empty + 8 * xs);
      const uint32_t sa = k & 1, pa = (k >> 1) & 1;
      const long long tq0 = dbg ? clock64() : 0;
      ptx::mbar_wait(a1_empty + 8 * sa, pa ^ 1);
      if (dbg) tw += clock64() - tq0;
      uint8_t* row = sm + C::A1_OFF + sa * C::A1_STAGE + r * 128;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(row + ((j ^ (r & 7)) << 4)) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
      ptx::fence_proxy_async_smem();
      ptx::mbar_arrive(a1_full + 8 * sa);
    }
    if (dbg) { long long* d = p.dbg + blockIdx.x * 16; d[5] = tw; d[6] = clock64() - tb0; d[13] = tx; }
  } else if (warp == UT_TMA_WARP) {
    // ================================ x5 windows: one box load per item, UT_XS items ahead of the builders
    if (ptx::elect_one()) {
      uint32_t k = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
        const int txi = it % p.tiles_x, t = it / p.tiles_x;
        const int tyi = t % p.tiles_y, n = t / p.tiles_y;
        const uint32_t xs = k % UT_XS, xp = (k / UT_XS) & 1;
        ptx::mbar_wait(x_empty + 8 * xs, xp ^ 1);
        ptx::mbar_expect_tx(x_full + 8 * xs, UT_X_BYTES);
        ptx::tma_load_4d(sbase + C::X_OFF + xs * UT_X_STAGE, &xmap, x_full + 8 * xs, 0, 15 * txi - 2, 3 * tyi - 2, n);
      }
    }
  } else if (warp == UT_MMA_WARP || warp == 3) {
    // ================================ one issuing thread per half h (warp 7: phases 0-1, warp 3: phases 2-3).  Per item k:
    // GEMM #1 of (k, h) into acc1[h], then GEMM #2 of (k-1, h) from the A2 planes of that half into Y[h].
    const int h = warp == UT_MMA_WARP ? 0 : 1;
    if (ptx::elect_one()) {
      if (h == 0) {
        ptx::mbar_expect_tx(w_bar, 256 * 128 + 4096);
        ptx::bulk_load(sbase + C::W1_OFF, p.w1pack, 256 * 128, w_bar);
        ptx::bulk_load(sbase + C::W2_OFF, p.w2pack, 4096, w_bar);
      }
      ptx::mbar_wait(w_bar, 0);
      constexpr uint32_t idesc1 = ptx::idesc_bf16_f32(128, 128);
      constexpr uint32_t idesc2 = ptx::idesc_bf16_f32(128, 32);
      const uint64_t a1d0 = ptx::smem_desc_sw128(sbase + C::A1_OFF, 0);
      const uint64_t w1d = ptx::smem_desc_sw128(sbase + C::W1_OFF + h * 16384, 0);
      const uint64_t w2d = ptx::smem_desc_sw128(sbase + C::W2_OFF, 0);
      const bool dbg = p.dbg != nullptr && h == 0;
      long long t1 = 0, t2 = 0, t3 = 0, t4 = 0, tq = 0;
      const long long tm0 = dbg ? clock64() : 0;
      auto gemm2 = [&](uint32_t j) {                  // unit (j, h): two phase planes, each [128 x 64] x [64 x 32]
        if (dbg) tq = clock64();
        ptx::mbar_wait(y_empty + 8 * h, (j & 1) ^ 1);  // the epilogue of item j-1 has drained this half of Y
        if (dbg) { const long long n = clock64(); t3 += n - tq; tq = n; }
        ptx::mbar_wait(a2_full + 8 * h, j & 1);
        if (dbg) t4 += clock64() - tq;
        ptx::tc_fence_after();
#pragma unroll
        for (int pp = 0; pp < 2; ++pp)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            mma_bf16_ts(tmem + TM_Y + (2 * h + pp) * 32, tmem + TM_A2 + (2 * h + pp) * 32 + kk * 8,
                        ptx::desc_advance(w2d, kk * 32), idesc2, kk != 0);
        ptx::tc_commit(a2_empty + 8 * h);
        ptx::tc_commit(y_full + 8 * (2 * (j % UT_EPI_GROUPS) + h));
      };
      uint32_t k = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
        const uint32_t sa = k & 1, pa = (k >> 1) & 1;
        if (dbg) tq = clock64();
        ptx::mbar_wait(a1_full + 8 * sa, pa);
        if (dbg) { const long long n = clock64(); t1 += n - tq; tq = n; }
        ptx::mbar_wait(acc1_empty + 8 * h, (k & 1) ^ 1);
        if (dbg) t2 += clock64() - tq;
        ptx::tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
          ptx::mma_bf16_ss(tmem + h * 128, ptx::desc_advance(a1d0, sa * C::A1_STAGE + kk * 32), ptx::desc_advance(w1d, kk * 32),
                           idesc1, kk);
        ptx::tc_commit(acc1_full + 8 * h);
        ptx::tc_commit(a1_empty + 8 * sa);             // the stage is free when BOTH halves' MMAs have read it (count 2)
        if (k > 0) gemm2(k - 1);
      }
      if (k > 0) gemm2(k - 1);
      if (dbg) {
        long long* d = p.dbg + blockIdx.x * 16;
        d[0] = clock64() - tm0; d[1] = t1; d[2] = t2; d[3] = t3; d[4] = t4;
      }
    }
  } else if (warp < 11) {   // warps 4-6 and 8-10 (3 and 7 issue the MMAs)
    // ================================ mid-drain: relu(upconv1) of low-res pixel r, phases 2h and 2h+1 (group h = warps 4-6 /
    // 8-10), back into TMEM as the bf16 A operand of the second GEMM - same lane, no shared memory.
    const int h = warp >> 3;
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int lr = r / UT_LC, lc = r - lr * UT_LC;
    const uint32_t tl = tmem + ((uint32_t)(q * 32) << 16);
    uint32_t k = 0;
    const bool dbg = p.dbg != nullptr && warp == 4 && lane == 0;
    long long t7 = 0, t8 = 0, tq = 0, tld = 0, tst = 0;
    const long long td0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
      const int txi = it % p.tiles_x, t = it / p.tiles_x;
      const int tyi = t % p.tiles_y;
      if (dbg) tq = clock64();
      ptx::mbar_wait(acc1_full + 8 * h, k & 1);
      if (dbg) t8 += clock64() - tq;
      ptx::tc_fence_after();
#pragma unroll
      for (int pp = 0; pp < 2; ++pp) {
        // hi-res pixel of this phase: halo-tile coordinates hr = 2*lr + py - 1, hc = 2*lc + px - 1 (py = h, px = pp)
        const int hr = 2 * lr + h - 1, hc = 2 * lc + pp - 1;
        const int hy = 6 * tyi - 1 + hr, hx = 30 * txi - 1 + hc;
        // rows outside the image are upconv2's zero padding; rows of the patch outside the halo tile are never read
        const bool keep = r < UT_ROWS1 && hy >= 0 && hy < H2 && hx >= 0 && hx < W2;
        uint32_t o[32];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          const long long tl0 = dbg ? clock64() : 0;
          ptx::tmem_ld_32x32(tl + h * 128 + pp * 64 + half * 32, v);
          ptx::tmem_ld_wait();
          if (dbg) tld += clock64() - tl0;
#pragma unroll
          for (int j = 0; j < 16; ++j)
            o[half * 16 + j] = keep ? ptx::pack_relu_bf16x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])) : 0u;
        }
        if (pp == 0) {
          if (dbg) tq = clock64();
          ptx::mbar_wait(a2_empty + 8 * h, (k & 1) ^ 1);   // GEMM #2 of the previous item has consumed these planes
          if (dbg) t7 += clock64() - tq;
          ptx::tc_fence_after();
        } else {                                           // the accumulator half has been read completely: hand it back
          ptx::tc_fence_before();                          // before the last store, GEMM #1 of the next item can start
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(acc1_empty + 8 * h);
        }
        const long long ts0 = dbg ? clock64() : 0;
        ptx::tmem_st_32x32(tl + TM_A2 + (2 * h + pp) * 32, o);
        if (pp == 1) ptx::tmem_st_wait();
        if (dbg) tst += clock64() - ts0;
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(a2_full + 8 * h);    // both A2 planes of this half are in tensor memory
    }
    if (dbg) { long long* d = p.dbg + blockIdx.x * 16; d[7] = t7; d[8] = t8; d[9] = clock64() - td0; d[14] = tld; d[15] = tst; }
  } else if (warp >= UT_EPI_WARP0) {
    // ================================ epilogue: Y planes -> Ys, 9-tap gather over the phase planes, clamp, uint8 rows
    const int grp = (warp - UT_EPI_WARP0) >> 2;
    float* Ys = reinterpret_cast<float*>(sm + C::YS_OFF + grp * C::YS_BYTES);
    uint8_t* ob = sm + C::OB_OFF + grp * C::OB_BYTES;
    const int q = warp & 3;
    const int et = ((warp - UT_EPI_WARP0) & 3) * 32 + lane;
    const float b0 = p.bias2[0], b1 = p.bias2[1], b2 = p.bias2[2];
    const int r = q * 32 + lane;                     // Y row (low-res pixel of the patch) this lane drains
    const int rlr = r / UT_LC, rlc = r - rlr * UT_LC;
    const bool dbg = p.dbg != nullptr && warp == UT_EPI_WARP0 && lane == 0;
    long long t10 = 0, n_items = 0;
    const long long te0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x + grp * gridDim.x, k = grp; it < total; it += UT_EPI_GROUPS * gridDim.x, k += UT_EPI_GROUPS) {
      const int txi = it % p.tiles_x, t = it / p.tiles_x;
      const int tyi = t % p.tiles_y, n = t / p.tiles_y;
      const int x0 = txi * G27_TW, y0 = tyi * G27_ROWS;
      const long long tq = dbg ? clock64() : 0;
      ptx::mbar_wait(y_full + 8 * (2 * grp), (k / UT_EPI_GROUPS) & 1);
      ptx::mbar_wait(y_full + 8 * (2 * grp + 1), (k / UT_EPI_GROUPS) & 1);
      if (dbg) { t10 += clock64() - tq; ++n_items; }
      ptx::tc_fence_after();
      if (q < 3) {                                   // lanes 96-127 hold no patch pixel
#pragma unroll
        for (int ph = 0; ph < 4; ++ph) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + TM_Y + ph * 32, v);
          ptx::tmem_ld_wait();
          const int hr = 2 * rlr + (ph >> 1) - 1, hc = 2 * rlc + (ph & 1) - 1;
          if (r < UT_ROWS1 && hr >= 0 && hr < 8 && hc >= 0 && hc < 32) {      // inside the 8 x 32 halo tile: the gather reads it
#pragma unroll
            for (int kk = 0; kk < 27; ++kk) Ys[kk * UT2_YS_PITCH + ph * UT2_PLANE + r] = __uint_as_float(v[kk]);
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) { ptx::mbar_arrive(y_empty); ptx::mbar_arrive(y_empty + 8); }
      }
      named_bar_sync(1 + grp, 128);
      const int nvx = min(G27_TW, W2 - x0);
#pragma unroll 1
      // one warp = one tile row (lanes 30, 31 idle)
      for (int idx = et; idx < G27_ROWS * 32; idx += 128) {
        const int ty = idx >> 5, tx = idx & 31;
        if (tx >= G27_TW) continue;
        const int y = y0 + ty, x = x0 + tx;
        float s0 = b0, s1 = b1, s2 = b2;
#pragma unroll
        for (int tp = 0; tp < 9; ++tp) {
          const int hr = ty + tp / 3, hc = tx + tp % 3;              // halo-tile coordinates of the tap's hi-res pixel
          const int m = (((hr + 1) & 1) * 2 + ((hc + 1) & 1)) * UT2_PLANE + ((hr + 1) >> 1) * UT_LC + ((hc + 1) >> 1);
          s0 += Ys[(tp * 3 + 0) * UT2_YS_PITCH + m];
          s1 += Ys[(tp * 3 + 1) * UT2_YS_PITCH + m];
          s2 += Ys[(tp * 3 + 2) * UT2_YS_PITCH + m];
        }
        if (y < H2 && x < W2) {
          const float c0 = fminf(fmaxf(s0, 0.f), 1.f), c1 = fminf(fmaxf(s1, 0.f), 1.f), c2 = fminf(fmaxf(s2, 0.f), 1.f);
          if (p.out_nchw) {
            const size_t plane = (size_t)H2 * W2;
            float* o = p.out_nchw + (size_t)n * 3 * plane + (size_t)y * W2 + x;
            o[0] = c0; o[plane] = c1; o[2 * plane] = c2;
          }
          if (p.out_u8) {
            // np.clip(x*255, 0, 255).astype(np.uint8): fp32 product, truncation (app.py:549)
            const uint8_t q0 = (uint8_t)__float2uint_rz(c0 * 255.0f), q1 = (uint8_t)__float2uint_rz(c1 * 255.0f),
                          q2 = (uint8_t)__float2uint_rz(c2 * 255.0f);
            const size_t rowb = (((size_t)n * H2 + y) * W2 + x0) * 3;
            const int ph = (int)(reinterpret_cast<uintptr_t>(p.out_u8 + rowb) & 3);
            uint8_t* sp = ob + ty * G27_OB_PITCH + ph + tx * 3;
            sp[0] = p.out_bgr ? q2 : q0; sp[1] = q1; sp[2] = p.out_bgr ? q0 : q2;
          }
        }
      }
      if (p.out_u8) {
        named_bar_sync(1 + grp, 128);
        for (int wi = et; wi < G27_ROWS * 24; wi += 128) {
          const int ty = wi / 24, j = wi - ty * 24;
          const int y = y0 + ty;
          if (y >= H2) continue;
          const size_t rowb = (((size_t)n * H2 + y) * W2 + x0) * 3;
          const int ph = (int)(reinterpret_cast<uintptr_t>(p.out_u8 + rowb) & 3), nbytes = nvx * 3;
          const int lo = j * 4, hi = lo + 4;
          if (lo >= ph + nbytes || hi <= ph) continue;
          uint8_t* gp = p.out_u8 + (rowb - ph) + lo;
          const uint8_t* sp = ob + ty * G27_OB_PITCH + lo;
          if (lo >= ph && hi <= ph + nbytes) {
            *reinterpret_cast<uint32_t*>(gp) = *reinterpret_cast<const uint32_t*>(sp);
          } else {
            for (int e = 0; e < 4; ++e)
              if (lo + e >= ph && lo + e < ph + nbytes) gp[e] = sp[e];
          }
        }
      }
      named_bar_sync(1 + grp, 128);
    }
    if (dbg) { long long* d = p.dbg + blockIdx.x * 16; d[10] = t10; d[11] = clock64() - te0; d[12] = n_items; }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == UT_MMA_WARP) {
    __syncwarp();
    ptx::tmem_dealloc(tmem, 512);
  }
}

}  // namespace thin
}  // namespace vex
