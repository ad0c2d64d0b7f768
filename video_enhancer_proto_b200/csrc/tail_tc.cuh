// Fused tail of AdvancedUpscaler: x5 -> upconv1 + relu -> upconv2 -> clamp -> uint8 (app.py:298-299, 549) in ONE
// kernel.  The `up` tensor ([n][2H][2W][64] bf16, 1.06 GB per 1080p frame - the largest activation of the network)
// never goes to HBM: unfused, writing and re-reading it costs more time than everything else in the two layers.
//
// Per item = one hi-res output tile of 6 rows x 30 px (halo tile 8 x 32 hi-res px = 5 x 17 low-res px):
//   x5 window  one TMA box load per item brings the 7 x 19 low-res pixels (fp32 x 4) under the patch and its 3x3 halo into
//              shared memory, 4 stages deep (out-of-image pixels arrive as zeros = the padding).  The first version gathered
//              the 9 neighbours per thread with global loads, one item ahead: the per-role cycle counters (option conv_dbg)
//              showed the MMA thread waiting 1255 of 2201 cycles per item for these builders - memory latency, not bandwidth.
//   builders   read the 3x3x3 neighbourhood of each of the 85 low-res pixels from the window (K = 27 -> 32), one
//              SWIZZLE_128B operand row per thread                                        (as im2col27, MODE_UP1)
//   MMA #1     [128 x 32] x [32 x 256]: ConvTranspose2d(3,64,4,2,1) as conv3x3 3 -> 4 phases x 64 ch  -> TMEM
//   mid-drain  TMEM -> bias, relu, bf16 -> PixelShuffle into the [256 hi-res px x 64 ch] A operand of the second
//              GEMM, in shared memory; hi-res pixels outside the image are zero rows (upconv2's zero padding)
//   MMA #2     2 x [128 x 64] x [64 x 32]: Y[p][(tap,co)], the 1x1-GEMM form of the 3x3 conv 64 -> 3   -> TMEM
//   epilogue   9-tap shifted gather-sum, clamp, x255, truncate, staged row stores           (as gemm27, MODE_UP2)
#pragma once

#include "thin_tc.cuh"

namespace vex {
namespace thin {

constexpr int UT_EPI_GROUPS = 3;       // the epilogue is the longest role: three groups of 4 warps take items k % 3
constexpr int UT_THREADS = 384 + 128 * UT_EPI_GROUPS;   // warps 0-3 builders; 4-6 / 8-10 mid-drain groups (phases 0-1 / 2-3;
                                       // TMEM lanes 0-95); 7 MMA; 11 TMA producer of the x5 windows; 12-15 / 16-19 / 20-23 epilogue groups

constexpr int UT_MMA_WARP = 7;
constexpr int UT_EPI_WARP0 = 12;
constexpr int UT_LR = 5, UT_LC = 17;   // low-res pixels under the 8 x 32 hi-res halo tile
constexpr int UT_ROWS1 = UT_LR * UT_LC;   // 85 live rows of the first GEMM
constexpr int UT_TMA_WARP = 11;
constexpr int UT_XS = 4;               // x5 window stages
constexpr int UT_XW = UT_LC + 2, UT_XH = UT_LR + 2;      // window = patch + 1-pixel halo: 19 x 7 low-res pixels
constexpr int UT_X_BYTES = UT_XW * UT_XH * 16;           // 2128: what one TMA box load delivers
constexpr int UT_X_STAGE = (UT_X_BYTES + 127) / 128 * 128;

struct UTCfg {
  static constexpr int A1_STAGE = 128 * 128;
  static constexpr int A1_OFF = 0;                                  // 2 stages
  static constexpr int W1_OFF = A1_OFF + 2 * A1_STAGE;              // 256 rows x 128 B
  static constexpr int A2_STAGE = 256 * 128;
  static constexpr int A2_OFF = W1_OFF + 256 * 128;                 // 2 stages
  static constexpr int W2_OFF = A2_OFF + 2 * A2_STAGE;              // 32 rows x 128 B
  static constexpr int YS_OFF = W2_OFF + 4096;
  static constexpr int YS_BYTES = 27 * G27_YS_PITCH * 4;
  static constexpr int OB_OFF = YS_OFF + UT_EPI_GROUPS * YS_BYTES;
  static constexpr int OB_BYTES = G27_ROWS * G27_OB_PITCH + 16;
  static constexpr int X_OFF = (OB_OFF + UT_EPI_GROUPS * OB_BYTES + 127) / 128 * 128;
  static constexpr int BAR_OFF = X_OFF + UT_XS * UT_X_STAGE;
  static constexpr int SMEM = BAR_OFF + 256 + 1024;
};

struct UTParams {
  const float4* x5;        // [n][H][W] conv5 output (3 channels + pad)
  const uint8_t* w1pack;   // 256 rows x 128 B SWIZZLE_128B image: row = phase*64 + co, k = (ky*3+kx)*3+ci for k < 27;
                           // k = 27, 28 hold the bias split into two bf16 (hi + lo) against constant-1 columns of A
  const uint8_t* w2pack;   // 32 rows x 128 B: row = tap*3+co, k = up channel
  const float* bias2;      // [3]
  uint8_t* out_u8;         // [n][2H][2W][3]
  float* out_nchw;         // optional [n][3][2H][2W]: clamp(...) before quantisation
  int H, W;                // LOW-res size
  int tiles_x, tiles_y, n_img;
  int out_bgr;
  float rcp_tx, rcp_ty;    // 1 / tiles_x, 1 / tiles_y when the item count is below 2^22 (else 0): tail_tc3.cuh divides by multiplying
  int probe;               // bring-up timing probes (WRONG RESULTS): 4 = no gather loads, 2 = mid-drain without stores, 1 = epilogue without gather
  long long* dbg;          // optional [gridDim.x][16] cycle counters per role (option conv_dbg): MMA thread 0 total, 1 wait a1_full,
                           // 2 wait acc1_empty, 3 wait acc2_empty, 4 wait a2_full; builder 5 wait a1_empty, 6 total; mid-drain (warp 4)
                           // 7 wait a2_empty, 8 wait acc1_full, 9 total; epilogue group 0: 10 wait acc2_full, 11 total, 12 items;
                           // 13 builder wait x_full, 14 mid-drain tcgen05.ld + wait
};

__global__ void __launch_bounds__(UT_THREADS, 1)
uptail_kernel(const __grid_constant__ CUtensorMap xmap, const UTParams p) {
  using C = UTCfg;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + C::BAR_OFF;
  const uint32_t a1_full = bar0, a1_empty = bar0 + 16;
  const uint32_t acc1_full = bar0 + 32, acc1_empty = bar0 + 48;      // per half of the first accumulator (phases 0-1 / 2-3)
  const uint32_t a2_full = bar0 + 64, a2_empty = bar0 + 80;
  const uint32_t acc2_full = bar0 + 96, acc2_empty = bar0 + 128;     // per epilogue group (up to 4)
  const uint32_t w_bar = bar0 + 160, tmem_slot = bar0 + 168;
  const uint32_t x_full = bar0 + 176, x_empty = bar0 + 176 + 8 * UT_XS;   // x5 window stages
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + C::BAR_OFF + 168);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = p.tiles_x * p.tiles_y * p.n_img;
  const int H2 = 2 * p.H, W2 = 2 * p.W;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(a1_full + 8 * i, 128); ptx::mbar_init(a1_empty + 8 * i, 1);
      ptx::mbar_init(a2_full + 8 * i, 192); ptx::mbar_init(a2_empty + 8 * i, 1);
      ptx::mbar_init(acc1_full + 8 * i, 1); ptx::mbar_init(acc1_empty + 8 * i, 3);
    }
    for (int i = 0; i < UT_EPI_GROUPS; ++i) { ptx::mbar_init(acc2_full + 8 * i, 1); ptx::mbar_init(acc2_empty + 8 * i, 4); }
    ptx::mbar_init(w_bar, 1);
    for (int i = 0; i < UT_XS; ++i) { ptx::mbar_init(x_full + 8 * i, 1); ptx::mbar_init(x_empty + 8 * i, 128); }
    ptx::fence_mbar_init();
    ptx::prefetch_tensormap(&xmap);
  }
  if (warp == UT_MMA_WARP) {
    ptx::tmem_alloc(tmem_slot, 512);                 // [0,256): first GEMM, two halves; 256 + grp*64 + g*32: second GEMM
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;

  if (warp < 4) {
    // ================================ builders: low-res pixel r of the 5 x 17 patch -> one operand row
    const int r = threadIdx.x;
    const int lr = r / UT_LC, lc = r - lr * UT_LC;
    uint32_t k = 0;
    const bool dbg = p.dbg != nullptr && threadIdx.x == 0;
    long long tw = 0, tx = 0, tb0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
      const int txi = it % p.tiles_x, t = it / p.tiles_x;
      const int tyi = t % p.tiles_y;
      const int y = 3 * tyi - 1 + lr, x = 15 * txi - 1 + lc;
      const bool live = r < UT_ROWS1 && y >= 0 && y < p.H && x >= 0 && x < p.W && !(p.probe & 4);
      const uint32_t xs = k % UT_XS, xp = (k / UT_XS) & 1;
      uint32_t w[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) w[j] = 0u;
      const long long tx0 = dbg ? clock64() : 0;
      ptx::mbar_wait(x_full + 8 * xs, xp);             // the window of this item has landed (TMA, 4 items ahead)
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
      ptx::mbar_arrive(x_empty + 8 * xs);              // the packed row depends on every window load: the stage may be refilled
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
    if (dbg) p.dbg[blockIdx.x * 16 + 13] = tx;
    if (dbg) { p.dbg[blockIdx.x * 16 + 5] = tw; p.dbg[blockIdx.x * 16 + 6] = clock64() - tb0; }
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
  } else if (warp == UT_MMA_WARP) {
    // ================================ both GEMMs from one thread; GEMM #2 of item k-1 is issued after GEMM #1 of item k
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(w_bar, 256 * 128 + 4096);
      ptx::bulk_load(sbase + C::W1_OFF, p.w1pack, 256 * 128, w_bar);
      ptx::bulk_load(sbase + C::W2_OFF, p.w2pack, 4096, w_bar);
      ptx::mbar_wait(w_bar, 0);
      constexpr uint32_t idesc1 = ptx::idesc_bf16_f32(128, 128);
      constexpr uint32_t idesc2 = ptx::idesc_bf16_f32(128, 32);
      // descriptors are encoded once; inside the loop an operand is one add away (ptx::desc_advance) - this thread's
      // serial instruction stream paces the whole kernel
      const uint64_t a1d0 = ptx::smem_desc_sw128(sbase + C::A1_OFF, 0), w1d = ptx::smem_desc_sw128(sbase + C::W1_OFF, 0);
      const uint64_t a2d0 = ptx::smem_desc_sw128(sbase + C::A2_OFF, 0), w2d = ptx::smem_desc_sw128(sbase + C::W2_OFF, 0);
      const bool dbg = p.dbg != nullptr;
      long long t1 = 0, t2 = 0, t3 = 0, t4 = 0, tq;
      const long long tm0 = dbg ? clock64() : 0;
      auto gemm2 = [&](uint32_t j) {
        const uint32_t s = j & 1, ph = (j >> 1) & 1;
        const uint32_t eg = j % UT_EPI_GROUPS, eu = j / UT_EPI_GROUPS;    // epilogue group and its use count
        if (dbg) tq = clock64();
        ptx::mbar_wait(acc2_empty + 8 * eg, (eu & 1) ^ 1);
        if (dbg) { const long long n = clock64(); t3 += n - tq; tq = n; }
        ptx::mbar_wait(a2_full + 8 * s, ph);
        if (dbg) t4 += clock64() - tq;
        ptx::tc_fence_after();
#pragma unroll
        for (int g = 0; g < 2; ++g)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const uint64_t ad = ptx::desc_advance(a2d0, s * C::A2_STAGE + g * 16384 + kk * 32);
            const uint64_t bd = ptx::desc_advance(w2d, kk * 32);
            ptx::mma_bf16_ss(tmem + 256 + eg * 64 + g * 32, ad, bd, idesc2, kk != 0);
          }
        ptx::tc_commit(a2_empty + 8 * s);
        ptx::tc_commit(acc2_full + 8 * eg);
      };
      uint32_t k = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
        const uint32_t sa = k & 1, pa = (k >> 1) & 1;
        if (dbg) tq = clock64();
        ptx::mbar_wait(a1_full + 8 * sa, pa);
        if (dbg) t1 += clock64() - tq;
#pragma unroll
        for (int h = 0; h < 2; ++h) {                // phases 0-1, then 2-3: the mid-drain of one half overlaps the other
          if (dbg) tq = clock64();
          ptx::mbar_wait(acc1_empty + 8 * h, (k & 1) ^ 1);
          if (dbg) t2 += clock64() - tq;
          ptx::tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            const uint64_t ad = ptx::desc_advance(a1d0, sa * C::A1_STAGE + kk * 32);
            const uint64_t bd = ptx::desc_advance(w1d, h * 16384 + kk * 32);
            ptx::mma_bf16_ss(tmem + h * 128, ad, bd, idesc1, kk);
          }
          ptx::tc_commit(acc1_full + 8 * h);
        }
        ptx::tc_commit(a1_empty + 8 * sa);
        if (k > 0) gemm2(k - 1);
      }
      if (k > 0) gemm2(k - 1);
      if (dbg) {
        long long* d = p.dbg + blockIdx.x * 16;
        d[0] = clock64() - tm0; d[1] = t1; d[2] = t2; d[3] = t3; d[4] = t4;
      }
    }
  } else if (warp < 11) {   // (warp 7 = MMA and warp 11 = TMA were taken above)
    // ================================ mid-drain: relu(upconv1) of low-res pixel r -> rows of A2.  Group h (warps 4-6 /
    // 8-10) takes phases 2h and 2h+1 = columns [128h, 128h+128) of the first accumulator.
    const int h = warp >> 3;
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int lr = r / UT_LC, lc = r - lr * UT_LC;
    uint32_t k = 0;
    const bool dbg = p.dbg != nullptr && warp == 4 && lane == 0;
    long long t7 = 0, t8 = 0, tq = 0, tld = 0;
    const long long td0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
      const int txi = it % p.tiles_x, t = it / p.tiles_x;
      const int tyi = t % p.tiles_y;
      const uint32_t s = k & 1;
      if (dbg) tq = clock64();
      ptx::mbar_wait(a2_empty + 8 * s, ((k >> 1) & 1) ^ 1);
      if (dbg) { const long long n = clock64(); t7 += n - tq; tq = n; }
      ptx::mbar_wait(acc1_full + 8 * h, k & 1);
      if (dbg) t8 += clock64() - tq;
      ptx::tc_fence_after();
      uint8_t* a2 = sm + C::A2_OFF + s * C::A2_STAGE;
#pragma unroll
      for (int pp = 0; pp < 2; ++pp) {
        // hi-res pixel of this phase in halo-tile coordinates: hr = 2*lr + py - 1, hc = 2*lc + px - 1 (py = h, px = pp)
        const int hr = 2 * lr + h - 1, hc = 2 * lc + pp - 1;
        const bool in_tile = r < UT_ROWS1 && hr >= 0 && hr < 8 && hc >= 0 && hc < 32;
        const int hy = 6 * tyi - 1 + hr, hx = 30 * txi - 1 + hc;
        const bool in_img = hy >= 0 && hy < H2 && hx >= 0 && hx < W2;
        // A2 row order: within a hi-res row the even columns come first, then the odd ones.  One PixelShuffle phase has
        // one column parity, so a warp's 16-byte stores then cover all 8 swizzle positions (4 wavefronts instead of 8).
        const int m = hr * 32 + (hc >> 1) + 16 * (hc & 1);
        uint8_t* row = a2 + (m >> 7) * 16384 + (m & 127) * 128;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          const long long tl0 = dbg ? clock64() : 0;
          ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + h * 128 + pp * 64 + half * 32, v);
          ptx::tmem_ld_wait();
          if (dbg) tld += clock64() - tl0;
          if (in_tile && !(p.probe & 2)) {
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              uint4 o = make_uint4(0u, 0u, 0u, 0u);
              if (in_img)
                o = make_uint4(ptx::pack_relu_bf16x2(__uint_as_float(v[j4 * 8 + 0]), __uint_as_float(v[j4 * 8 + 1])),
                               ptx::pack_relu_bf16x2(__uint_as_float(v[j4 * 8 + 2]), __uint_as_float(v[j4 * 8 + 3])),
                               ptx::pack_relu_bf16x2(__uint_as_float(v[j4 * 8 + 4]), __uint_as_float(v[j4 * 8 + 5])),
                               ptx::pack_relu_bf16x2(__uint_as_float(v[j4 * 8 + 6]), __uint_as_float(v[j4 * 8 + 7])));
              *reinterpret_cast<uint4*>(row + (((half * 4 + j4) ^ (m & 7)) << 4)) = o;
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc1_empty + 8 * h);   // this half of the first accumulator is free
      ptx::fence_proxy_async_smem();
      ptx::mbar_arrive(a2_full + 8 * s);
    }
    if (dbg) { long long* d = p.dbg + blockIdx.x * 16; d[7] = t7; d[8] = t8; d[9] = clock64() - td0; d[14] = tld; }
  } else if (warp >= UT_EPI_WARP0) {
    // ================================ epilogue: as gemm27 MODE_UP2, on the hi-res tile
    const int grp = (warp - UT_EPI_WARP0) >> 2;
    float* Ys = reinterpret_cast<float*>(sm + C::YS_OFF + grp * C::YS_BYTES);
    uint8_t* ob = sm + C::OB_OFF + grp * C::OB_BYTES;
    const int q = warp & 3;
    const int et = ((warp - UT_EPI_WARP0) & 3) * 32 + lane;
    const float b0 = p.bias2[0], b1 = p.bias2[1], b2 = p.bias2[2];
    const uint32_t b = grp;
    uint32_t pb = 0;
    const bool dbg = p.dbg != nullptr && warp == UT_EPI_WARP0 && lane == 0;
    long long t10 = 0, n_items = 0;
    const long long te0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x + grp * gridDim.x; it < total; it += UT_EPI_GROUPS * gridDim.x) {
      const int txi = it % p.tiles_x, t = it / p.tiles_x;
      const int tyi = t % p.tiles_y, n = t / p.tiles_y;
      const int x0 = txi * G27_TW, y0 = tyi * G27_ROWS;
      const long long tq = dbg ? clock64() : 0;
      ptx::mbar_wait(acc2_full + 8 * b, pb);
      if (dbg) { t10 += clock64() - tq; ++n_items; }
      ptx::tc_fence_after();
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + 256 + b * 64 + g * 32, v);
        ptx::tmem_ld_wait();
        const int m = g * 128 + q * 32 + lane;
#pragma unroll
        for (int kk = 0; kk < 27; ++kk) Ys[kk * G27_YS_PITCH + m] = __uint_as_float(v[kk]);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc2_empty + 8 * b);
      pb ^= 1;
      named_bar_sync(1 + grp, 128);
      const int nvx = min(G27_TW, W2 - x0);
#pragma unroll 1
      // one warp = one tile row (lanes 30, 31 idle): a warp's 32 loads then fall into one row of Ys, i.e. 32 banks
      for (int idx = (p.probe & 1) ? 1000 : et; idx < G27_ROWS * 32; idx += 128) {
        const int ty = idx >> 5, tx = idx & 31;
        if (tx >= G27_TW) continue;
        const int y = y0 + ty, x = x0 + tx;
        float s0 = b0, s1 = b1, s2 = b2;
#pragma unroll
        for (int tp = 0; tp < 9; ++tp) {
          const int c = tx + tp % 3;                                  // halo-tile column; rows are stored even | odd
          const int m = (ty + tp / 3) * 32 + (c >> 1) + 16 * (c & 1);
          s0 += Ys[(tp * 3 + 0) * G27_YS_PITCH + m];
          s1 += Ys[(tp * 3 + 1) * G27_YS_PITCH + m];
          s2 += Ys[(tp * 3 + 2) * G27_YS_PITCH + m];
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
