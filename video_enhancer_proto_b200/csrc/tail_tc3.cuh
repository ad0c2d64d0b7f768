// Fused tail of AdvancedUpscaler, generation 4 (option fuse_tail = 4): tail_tc.cuh with the second GEMM and the epilogue
// rebuilt around one observation - the A operand of GEMM #2 already IS an 8-row x 32-column hi-res halo tile at a pitch of
// 32 pixels, so a warp's 32 TMEM lanes are the 32 columns of one tile row.
//
//   tail_tc.cuh   GEMM #2 = 2 x [128 x 64] x [64 x 32]: Y[p][(tap, co)], 27 planes per pixel; the epilogue copies them from
//                 TMEM to 27 shared-memory planes and gathers 9 taps x 3 channels per output through shared memory.  That
//                 epilogue is the longest role of the kernel (5900 clocks per item over three groups of four warps).
//   here          the VERTICAL tap is accumulated by the tensor core: D[m][dx*3 + co] = sum over dy, k of
//                 A2[m + 32*dy][k] * W2[dy][dx*3 + co][k] - three MMAs of N = 16 per K = 16 step, the A descriptor advanced
//                 by whole tile rows (4 KB, on the swizzle atom) - and the HORIZONTAL tap is three lane shuffles in the
//                 epilogue: out[x] = D[x-1][0:3] + D[x][3:6] + D[x+1][6:9].  No planes, no gather: the epilogue reads 9
//                 values per pixel from TMEM and one group of four warps is enough.  (Same construction as rrdb_tc.cuh.)
// 24 MMAs of N = 16 replace 8 of N = 32; one thread issuing them all would pace the kernel (~65 clocks of its own
// instruction stream per MMA), so GEMM #2 has two issuing threads of its own, one per accumulator (warps 12, 13), and warp 7
// keeps GEMM #1.  Summation order differs from tail_tc.cuh (fp32 in the tensor core over dy and k, then left + centre +
// right + bias), so the two tails agree to fp32 rounding, not bit for bit: the uint8 outputs differ by at most 1 LSB on a
// small fraction of bytes (tests/test_cnn_gpu.py).
//
// Everything up to the A2 operand (x5 windows by TMA, builders, GEMM #1, mid-drain) is tail_tc.cuh's.
#pragma once

#include <type_traits>

#include "tail_tc.cuh"

namespace vex {
namespace thin {

// Item = 10 x 30 hi-res outputs (halo tile 12 x 32 = 384 rows of the second GEMM, three accumulators) from a 7 x 17 low-res
// patch: 119 of the 128 rows of GEMM #1 are live (tail_tc.cuh: 5 x 17 = 85 rows for 6 x 30 outputs).  Every role pays its
// barrier round trips and its dependent per-lane chain once per item, so 1.67x the outputs per item is the lever that is
// left; tail_tc.cuh cannot take it because its 27 gather planes for 384 rows would not fit shared memory.
// 28 warps: 0-3 builders | 4-7, 8-11 mid-drain (phases 0-1 / 2-3; warp & 3 = TMEM lane quarter) | 12 GEMM #1 | 13 TMA |
// 14-16 GEMM #2 (one accumulator each) | 20-23, 24-27 epilogue groups (items k % 2).  A 32-warp layout with four mid-drain
// groups was measured on the small item: every role got SLOWER (2620 clocks per item against 1990).
constexpr int UT3_LR = 7, UT3_LC = 17;             // low-res pixels under the 12 x 32 hi-res halo tile
constexpr int UT3_ROWS1 = UT3_LR * UT3_LC;         // 119 live rows of the first GEMM
constexpr int UT3_OROWS = 10;                      // hi-res output rows per item
constexpr int UT3_HROWS = UT3_OROWS + 2;           // halo tile rows
constexpr int UT3_NACC2 = 3;                       // accumulators of GEMM #2 (4 tile rows each)
constexpr int UT3_XW = UT3_LC + 2, UT3_XH = UT3_LR + 2;   // x5 window: 19 x 9 low-res pixels
constexpr int UT3_X_BYTES = UT3_XW * UT3_XH * 16;
constexpr int UT3_X_STAGE = (UT3_X_BYTES + 127) / 128 * 128;
constexpr int UT3_EPI_GROUPS = 2;
constexpr int UT3_MMA1_WARP = 12, UT3_TMA_WARP = 13, UT3_MMA2_WARP0 = 14;
constexpr int UT3_EPI_WARP0 = 20;      // warp & 3 = TMEM lane quarter
constexpr int UT3_THREADS = 32 * (UT3_EPI_WARP0 + 4 * UT3_EPI_GROUPS);   // 896
constexpr int UT3_NACC1 = 3;           // ring of half-accumulators (128 columns each) of GEMM #1

struct UT3Cfg {
  static constexpr int A1_STAGE = 128 * 128;
  static constexpr int A1_OFF = 0;                                  // 2 stages
  static constexpr int W1_OFF = A1_OFF + 2 * A1_STAGE;              // 256 rows x 128 B
  static constexpr int A2_STAGE = UT3_HROWS * 32 * 128;
  static constexpr int A2_OFF = W1_OFF + 256 * 128;                 // 2 stages
  static constexpr int W2_BYTES = 3 * 16 * 128;                     // [dy] slices of 16 rows (dx*3 + co, 9 used) x 64 k
  static constexpr int W2_OFF = A2_OFF + 2 * A2_STAGE;
  static constexpr int PAD_OFF = W2_OFF + 8192;                     // GEMM #2 of the last accumulator, dy = 2 reads 8 KB past its A2 stage
  static constexpr int OB_OFF = PAD_OFF + 8192;                     // (rows of outputs that do not exist): keep that finite, zeroed
  static constexpr int OB_BYTES = UT3_OROWS * G27_OB_PITCH + 16;
  static constexpr int X_OFF = (OB_OFF + UT3_EPI_GROUPS * OB_BYTES + 127) / 128 * 128;
  static constexpr int BAR_OFF = X_OFF + UT_XS * UT3_X_STAGE;
  static constexpr int SMEM = BAR_OFF + 512 + 1024;
};

// it -> (txi, tyi, n).  Every warp of every role needs this once per item, and two 32-bit integer divisions are ~60
// instructions each time - a quarter of all warp-instructions the kernel issues per item, on schedulers that are 63 %
// busy.  For item counts below 2^22 the quotient is a float multiply plus one correction step (exact: the estimate is off by
// at most one).
__device__ __forceinline__ void ut3_tile(const UTParams& p, int it, int& txi, int& tyi, int& n) {
  if (p.rcp_tx != 0.f) {
    int t = __float2int_rz(__int2float_rn(it) * p.rcp_tx);
    int r = it - t * p.tiles_x;
    if (r < 0) { --t; r += p.tiles_x; } else if (r >= p.tiles_x) { ++t; r -= p.tiles_x; }
    txi = r;
    int q = __float2int_rz(__int2float_rn(t) * p.rcp_ty);
    int r2 = t - q * p.tiles_y;
    if (r2 < 0) { --q; r2 += p.tiles_y; } else if (r2 >= p.tiles_y) { ++q; r2 -= p.tiles_y; }
    tyi = r2; n = q;
  } else {
    txi = it % p.tiles_x;
    const int t = it / p.tiles_x;
    tyi = t % p.tiles_y; n = t / p.tiles_y;
  }
}

// DBG = the per-role cycle counters of option conv_dbg.  They are a template parameter, not a run-time flag: predicated-off
// instructions still take issue slots, and issue slots are what this kernel is short of.
template <bool DBG, bool NCHW>   // NCHW: the fp32 pre-quantisation output of the parity hook (p.out_nchw) is written too
__global__ void __launch_bounds__(UT3_THREADS, 1)
uptail3_kernel(const __grid_constant__ CUtensorMap xmap, const UTParams p) {
  using C = UT3Cfg;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + C::BAR_OFF;
  const uint32_t a1_full = bar0, a1_empty = bar0 + 16;
  // GEMM #1 works through half-jobs j = 2k + h (item k, phases 2h and 2h+1) on a RING of UT3_NACC1 half-accumulators: in
  // tail_tc.cuh a half has ONE buffer, so "mid-drain drains it (~1450 clocks, a latency-bound chain)" and "GEMM #1 refills
  // it (~400 clocks until the barrier fires)" take turns, and their sum is the ~1960 clocks per item every earlier variant
  // of this kernel measured.  With a third buffer the next item's first half is computed while this item's halves drain.
  const uint32_t acc1_full = bar0 + 256, acc1_empty = bar0 + 288;
  const uint32_t a2_full = bar0 + 64, a2_empty = bar0 + 80;
  const uint32_t acc2_full = bar0 + 96, acc2_empty = bar0 + 128;     // per epilogue group
  const uint32_t w_bar = bar0 + 160, tmem_slot = bar0 + 168;
  const uint32_t x_full = bar0 + 176, x_empty = bar0 + 176 + 8 * UT_XS;   // x5 window stages
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + C::BAR_OFF + 168);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = p.tiles_x * p.tiles_y * p.n_img;
  const int H2 = 2 * p.H, W2 = 2 * p.W;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      // producers arrive ONCE PER WARP (fence, __syncwarp, lane 0): 4 builder warps, 6 mid-drain warps.  tail_tc.cuh has every
      // thread arrive - ~450 shared-memory atomics per item on three barriers
      ptx::mbar_init(a1_full + 8 * i, 4); ptx::mbar_init(a1_empty + 8 * i, 1);
      ptx::mbar_init(a2_full + 8 * i, 8); ptx::mbar_init(a2_empty + 8 * i, UT3_NACC2);   // every GEMM #2 thread releases a stage
    }
    for (int i = 0; i < UT3_NACC1; ++i) { ptx::mbar_init(acc1_full + 8 * i, 1); ptx::mbar_init(acc1_empty + 8 * i, 4); }
    for (int i = 0; i < UT3_EPI_GROUPS; ++i) { ptx::mbar_init(acc2_full + 8 * i, UT3_NACC2); ptx::mbar_init(acc2_empty + 8 * i, 4); }
    ptx::mbar_init(w_bar, 1);
    for (int i = 0; i < UT_XS; ++i) { ptx::mbar_init(x_full + 8 * i, 1); ptx::mbar_init(x_empty + 8 * i, 4); }
    ptx::fence_mbar_init();
    ptx::prefetch_tensormap(&xmap);
  }
  for (int i = threadIdx.x; i < 8192 / 16; i += UT3_THREADS) reinterpret_cast<uint4*>(sm + C::PAD_OFF)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = threadIdx.x; i < 8192 / 16; i += UT3_THREADS) reinterpret_cast<uint4*>(sm + C::W2_OFF)[i] = make_uint4(0u, 0u, 0u, 0u);
  ptx::fence_proxy_async_smem();
  if (warp == UT3_MMA1_WARP) {
    ptx::tmem_alloc(tmem_slot, 512);                 // [0,384): first GEMM, three half-accumulators; 384 + grp*64 + g*16: second GEMM
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;

  if (warp < 4) {
    // ================================ builders: low-res pixel r of the 5 x 17 patch -> one operand row
    const int r = threadIdx.x;
    const int lr = r / UT3_LC, lc = r - lr * UT3_LC;
    uint32_t k = 0;
    const bool dbg = DBG && p.dbg != nullptr && threadIdx.x == 0;
    long long tw = 0, tx = 0, tb0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
      int txi, tyi, n_unused;
      ut3_tile(p, it, txi, tyi, n_unused);
      const int y = (UT3_OROWS / 2) * tyi - 1 + lr, x = 15 * txi - 1 + lc;
      const bool live = r < UT3_ROWS1 && y >= 0 && y < p.H && x >= 0 && x < p.W && !(p.probe & 4);
      const uint32_t xs = k % UT_XS, xp = (k / UT_XS) & 1;
      uint32_t w[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) w[j] = 0u;
      const long long tx0 = dbg ? clock64() : 0;
      ptx::mbar_wait(x_full + 8 * xs, xp);             // the window of this item has landed (TMA, 4 items ahead)
      if (dbg) tx += clock64() - tx0;
      if (live) {
        const float4* win = reinterpret_cast<const float4*>(sm + C::X_OFF + xs * UT3_X_STAGE) + lr * UT3_XW + lc;
        float v[27];
#pragma unroll
        for (int q9 = 0; q9 < 9; ++q9) {
          const float4 f = win[(q9 / 3) * UT3_XW + q9 % 3];
          v[q9 * 3 + 0] = f.x; v[q9 * 3 + 1] = f.y; v[q9 * 3 + 2] = f.z;
        }
#pragma unroll
        for (int j = 0; j < 13; ++j) w[j] = ptx::pack_bf16x2(v[2 * j], v[2 * j + 1]);
        w[13] = ptx::pack_bf16x2(v[26], 1.0f);       // k = 27, 28: constant 1 against the bias rows of W1
        w[14] = ptx::pack_bf16x2(1.0f, 0.f);
      }
      __syncwarp();                                    // the packed rows depend on every window load: the stage may be refilled
      if (lane == 0) ptx::mbar_arrive(x_empty + 8 * xs);
      const uint32_t sa = k & 1, pa = (k >> 1) & 1;
      const long long tq0 = dbg ? clock64() : 0;
      ptx::mbar_wait(a1_empty + 8 * sa, pa ^ 1);
      if (dbg) tw += clock64() - tq0;
      uint8_t* row = sm + C::A1_OFF + sa * C::A1_STAGE + r * 128;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(row + ((j ^ (r & 7)) << 4)) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(a1_full + 8 * sa);
    }
    if (dbg) p.dbg[blockIdx.x * 16 + 13] = tx;
    if (dbg) { p.dbg[blockIdx.x * 16 + 5] = tw; p.dbg[blockIdx.x * 16 + 6] = clock64() - tb0; }
  } else if (warp == UT3_TMA_WARP) {
    // ================================ x5 windows: one box load per item, UT_XS items ahead of the builders
    if (ptx::elect_one()) {
      uint32_t k = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
        int txi, tyi, n;
        ut3_tile(p, it, txi, tyi, n);
        const uint32_t xs = k % UT_XS, xp = (k / UT_XS) & 1;
        ptx::mbar_wait(x_empty + 8 * xs, xp ^ 1);
        ptx::mbar_expect_tx(x_full + 8 * xs, UT3_X_BYTES);
        ptx::tma_load_4d(sbase + C::X_OFF + xs * UT3_X_STAGE, &xmap, x_full + 8 * xs, 0, 15 * txi - 2, (UT3_OROWS / 2) * tyi - 2, n);
      }
    }
  } else if (warp == UT3_MMA1_WARP) {
    // ================================ GEMM #1 (and the weight loads)
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(w_bar, 256 * 128 + C::W2_BYTES);
      ptx::bulk_load(sbase + C::W1_OFF, p.w1pack, 256 * 128, w_bar);
      ptx::bulk_load(sbase + C::W2_OFF, p.w2pack, C::W2_BYTES, w_bar);
      ptx::mbar_wait(w_bar, 0);
      constexpr uint32_t idesc1 = ptx::idesc_bf16_f32(128, 128);
      const uint64_t a1d0 = ptx::smem_desc_sw128(sbase + C::A1_OFF, 0), w1d = ptx::smem_desc_sw128(sbase + C::W1_OFF, 0);
      const bool dbg = DBG && p.dbg != nullptr;
      long long t1 = 0, t2 = 0, tq;
      const long long tm0 = dbg ? clock64() : 0;
      uint32_t k = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
        const uint32_t sa = k & 1, pa = (k >> 1) & 1;
        if (dbg) tq = clock64();
        ptx::mbar_wait(a1_full + 8 * sa, pa);
        if (dbg) t1 += clock64() - tq;
#pragma unroll
        for (int h = 0; h < 2; ++h) {                // phases 0-1, then 2-3: the mid-drain of one half overlaps the other
          const uint32_t jb = (2 * k + h) % UT3_NACC1, ju = (2 * k + h) / UT3_NACC1;   // ring buffer and its use count
          if (dbg) tq = clock64();
          ptx::mbar_wait(acc1_empty + 8 * jb, (ju & 1) ^ 1);
          if (dbg) t2 += clock64() - tq;
          ptx::tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            const uint64_t ad = ptx::desc_advance(a1d0, sa * C::A1_STAGE + kk * 32);
            const uint64_t bd = ptx::desc_advance(w1d, h * 16384 + kk * 32);
            ptx::mma_bf16_ss(tmem + jb * 128, ad, bd, idesc1, kk);
          }
          ptx::tc_commit(acc1_full + 8 * jb);
        }
        ptx::tc_commit(a1_empty + 8 * sa);
      }
      if (dbg) {
        long long* d = p.dbg + blockIdx.x * 16;
        d[0] = clock64() - tm0; d[1] = t1; d[2] = t2;
      }
    }
  } else if (warp >= UT3_MMA2_WARP0 && warp < UT3_MMA2_WARP0 + UT3_NACC2) {
    // ================================ GEMM #2, accumulator g: output rows 4g .. 4g+3 of the tile (rows 6, 7 do not exist)
    const int g = warp - UT3_MMA2_WARP0;
    if (ptx::elect_one()) {
      ptx::mbar_wait(w_bar, 0);
      constexpr uint32_t idesc2 = ptx::idesc_bf16_f32(128, 16);
      const uint64_t a2d0 = ptx::smem_desc_sw128(sbase + C::A2_OFF + (4 * g) * 32 * 128, 0);
      const uint64_t w2d = ptx::smem_desc_sw128(sbase + C::W2_OFF, 0);
      const bool dbg = DBG && p.dbg != nullptr && g == 0;
      long long t3 = 0, t4 = 0, tq;
      uint32_t k = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
        const uint32_t s = k & 1, ph = (k >> 1) & 1;
        const uint32_t eg = k % UT3_EPI_GROUPS, eu = k / UT3_EPI_GROUPS;   // epilogue group and its use count
        if (dbg) tq = clock64();
        ptx::mbar_wait(acc2_empty + 8 * eg, (eu & 1) ^ 1);
        if (dbg) { const long long n = clock64(); t3 += n - tq; tq = n; }
        ptx::mbar_wait(a2_full + 8 * s, ph);
        if (dbg) t4 += clock64() - tq;
        ptx::tc_fence_after();
        const uint32_t dcol = tmem + 384 + eg * 64 + g * 16;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)               // output row r reads halo rows r, r+1, r+2
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const uint64_t ad = ptx::desc_advance(a2d0, s * C::A2_STAGE + dy * 32 * 128 + kk * 32);
            const uint64_t bd = ptx::desc_advance(w2d, dy * 2048 + kk * 32);
            ptx::mma_bf16_ss(dcol, ad, bd, idesc2, (dy | kk) != 0);
          }
        ptx::tc_commit(a2_empty + 8 * s);
        ptx::tc_commit(acc2_full + 8 * eg);
      }
      if (dbg) { long long* d = p.dbg + blockIdx.x * 16; d[3] = t3; d[4] = t4; }
    }
  } else if (warp < 12) {
    // ================================ mid-drain: relu(upconv1) of low-res pixel r -> rows of A2.  Group h (warps 4-7 /
    // 8-11) takes phases 2h and 2h+1 = one half-accumulator of 128 columns.
    const int h = (warp >> 2) - 1;
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int lr = r / UT3_LC, lc = r - lr * UT3_LC;
    uint32_t k = 0;
    const bool dbg = DBG && p.dbg != nullptr && warp == 4 && lane == 0;
    long long t7 = 0, t8 = 0, tq = 0, tld = 0;
    const long long td0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x; it < total; it += gridDim.x, ++k) {
      int txi, tyi, n_unused;
      ut3_tile(p, it, txi, tyi, n_unused);
      const uint32_t s = k & 1;
      if (dbg) tq = clock64();
      ptx::mbar_wait(a2_empty + 8 * s, ((k >> 1) & 1) ^ 1);
      if (dbg) { const long long n = clock64(); t7 += n - tq; tq = n; }
      const uint32_t jb = (2 * k + h) % UT3_NACC1, ju = (2 * k + h) / UT3_NACC1;
      ptx::mbar_wait(acc1_full + 8 * jb, ju & 1);
      if (dbg) t8 += clock64() - tq;
      ptx::tc_fence_after();
      uint8_t* a2 = sm + C::A2_OFF + s * C::A2_STAGE;
      // An item whose halo tile lies inside the image (all but the border items) needs no per-pixel image test and no
      // zero rows: the same loop is instantiated without them (fewer instructions on the role that paces the kernel).
      const bool interior = UT3_OROWS * tyi >= 1 && UT3_OROWS * tyi + UT3_HROWS - 1 <= H2 && 30 * txi >= 1 && 30 * txi + 31 <= W2;
      auto drain = [&](auto interior_c) {
      constexpr bool INTERIOR = decltype(interior_c)::value;
#pragma unroll
      for (int pp = 0; pp < 2; ++pp) {
        // hi-res pixel of this phase in halo-tile coordinates: hr = 2*lr + py - 1, hc = 2*lc + px - 1 (py = h, px = pp)
        const int hr = 2 * lr + h - 1, hc = 2 * lc + pp - 1;
        const bool in_tile = r < UT3_ROWS1 && hr >= 0 && hr < UT3_HROWS && hc >= 0 && hc < 32;
        const int hy = UT3_OROWS * tyi - 1 + hr, hx = 30 * txi - 1 + hc;
        const bool in_img = INTERIOR || (hy >= 0 && hy < H2 && hx >= 0 && hx < W2);
        // A2 row order: within a hi-res row the even columns come first, then the odd ones.  One PixelShuffle phase has
        // one column parity, so a warp's 16-byte stores then cover all 8 swizzle positions (4 wavefronts instead of 8).
        const int m = hr * 32 + (hc >> 1) + 16 * (hc & 1);
        uint8_t* row = a2 + m * 128;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          const long long tl0 = dbg ? clock64() : 0;
          ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + jb * 128 + pp * 64 + half * 32, v);
          ptx::tmem_ld_wait();
          if (dbg) tld += clock64() - tl0;
          if (in_tile) {
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
      };
      if (interior) drain(std::true_type{}); else drain(std::false_type{});
      ptx::tc_fence_before();
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive(acc1_empty + 8 * jb);               // this half-accumulator is free
        ptx::mbar_arrive(a2_full + 8 * s);
      }
    }
    if (dbg) { long long* d = p.dbg + blockIdx.x * 16; d[7] = t7; d[8] = t8; d[9] = clock64() - td0; d[14] = tld; }
  } else if (warp >= UT3_EPI_WARP0) {
    // ================================ epilogue: lane = position in the tile row (even columns 0-15, odd columns 16-31, the
    // order the mid-drain stores them in); warp quarter q = tile row 4g + q of accumulator g
    const int grp = (warp - UT3_EPI_WARP0) >> 2;
    uint8_t* ob = sm + C::OB_OFF + grp * C::OB_BYTES;
    const int q = warp & 3;
    const int et = ((warp - UT3_EPI_WARP0) & 3) * 32 + lane;
    const float b0 = p.bias2[0], b1 = p.bias2[1], b2 = p.bias2[2];
    const bool even = lane < 16;
    const int j = even ? lane : lane - 16;
    const int hc = even ? 2 * j : 2 * j + 1;            // halo-tile column of this lane; output column tx = hc - 1
    const int src_l = even ? 16 + ((j + 15) & 15) : j;  // lane holding column hc - 1 (column 0 has none: masked below)
    const int src_r = even ? 16 + j : (j + 1) & 15;     // lane holding column hc + 1 (column 31 has none)
    const int tx = hc - 1;
    uint32_t pb = 0;
    const bool dbg = DBG && p.dbg != nullptr && warp == UT3_EPI_WARP0 && lane == 0;
    long long t10 = 0, n_items = 0;
    const long long te0 = dbg ? clock64() : 0;
    for (int it = blockIdx.x + grp * gridDim.x; it < total; it += UT3_EPI_GROUPS * gridDim.x) {
      int txi, tyi, n;
      ut3_tile(p, it, txi, tyi, n);
      const int x0 = txi * G27_TW, y0 = tyi * UT3_OROWS;
      const long long tq = dbg ? clock64() : 0;
      ptx::mbar_wait(acc2_full + 8 * grp, pb);
      if (dbg) { t10 += clock64() - tq; ++n_items; }
      ptx::tc_fence_after();
      uint32_t va[16], vb[16], vc[16];
      ptx::tmem_ld_32x16(tmem + ((uint32_t)(q * 32) << 16) + 384 + grp * 64, va);        // accumulator 0: tile row q
      ptx::tmem_ld_32x16(tmem + ((uint32_t)(q * 32) << 16) + 384 + grp * 64 + 16, vb);   // accumulator 1: tile row 4 + q
      ptx::tmem_ld_32x16(tmem + ((uint32_t)(q * 32) << 16) + 384 + grp * 64 + 32, vc);   // accumulator 2: tile row 8 + q (rows 10, 11 do not exist)
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc2_empty + 8 * grp);
      pb ^= 1;
      const int nvx = min(G27_TW, W2 - x0);
      // low address bits of the first byte of output row y0 + q; a row further down adds 3 * W2 (only bits 0-1 matter)
      const uint32_t row_step_lo = (uint32_t)(3 * W2);
      const uint32_t row0_lo = (uint32_t)reinterpret_cast<uintptr_t>(p.out_u8) +
                               (uint32_t)(((uint32_t)n * (uint32_t)H2 + (uint32_t)(y0 + q)) * row_step_lo) + (uint32_t)(3 * x0);
#pragma unroll
      for (int g = 0; g < UT3_NACC2; ++g) {
        const uint32_t* v = g == 0 ? va : (g == 1 ? vb : vc);
        float s[3];
#pragma unroll
        for (int co = 0; co < 3; ++co) {
          const float left = __shfl_sync(0xffffffffu, __uint_as_float(v[co]), src_l);        // D[hc-1][dx = 0]
          const float right = __shfl_sync(0xffffffffu, __uint_as_float(v[6 + co]), src_r);   // D[hc+1][dx = 2]
          s[co] = (left + __uint_as_float(v[3 + co])) + right;
        }
        const int ty = 4 * g + q;
        const int y = y0 + ty, x = x0 + tx;
        if (ty < UT3_OROWS && tx >= 0 && tx < G27_TW && y < H2 && x < W2 && !(p.probe & 1)) {
          const float c0 = fminf(fmaxf(s[0] + b0, 0.f), 1.f), c1 = fminf(fmaxf(s[1] + b1, 0.f), 1.f),
                      c2 = fminf(fmaxf(s[2] + b2, 0.f), 1.f);
          if (NCHW && p.out_nchw) {
            const size_t plane = (size_t)H2 * W2;
            float* o = p.out_nchw + (size_t)n * 3 * plane + (size_t)y * W2 + x;
            o[0] = c0; o[plane] = c1; o[2 * plane] = c2;
          }
          if (p.out_u8) {
            // np.clip(x*255, 0, 255).astype(np.uint8): fp32 product, truncation (app.py:549)
            const uint8_t q0 = (uint8_t)__float2uint_rz(c0 * 255.0f), q1 = (uint8_t)__float2uint_rz(c1 * 255.0f),
                          q2 = (uint8_t)__float2uint_rz(c2 * 255.0f);
            const int ph = (int)((row0_lo + (uint32_t)(4 * g) * row_step_lo) & 3u);   // (out_u8 + row start of y) & 3
            uint8_t* sp = ob + ty * G27_OB_PITCH + ph + tx * 3;
            sp[0] = p.out_bgr ? q2 : q0; sp[1] = q1; sp[2] = p.out_bgr ? q0 : q2;
          }
        }
      }
      if (p.out_u8) {
        named_bar_sync(1 + grp, 128);
        for (int wi = et; wi < UT3_OROWS * 24; wi += 128) {
          const int ty = wi / 24, jw = wi - ty * 24;
          const int y = y0 + ty;
          if (y >= H2) continue;
          const size_t rowb = (((size_t)n * H2 + y) * W2 + x0) * 3;
          const int ph = (int)(reinterpret_cast<uintptr_t>(p.out_u8 + rowb) & 3), nbytes = nvx * 3;
          const int lo = jw * 4, hi = lo + 4;
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
        named_bar_sync(1 + grp, 128);
      }
    }
    if (dbg) { long long* d = p.dbg + blockIdx.x * 16; d[10] = t10; d[11] = clock64() - te0; d[12] = n_items; }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == UT3_MMA1_WARP) {
    __syncwarp();
    ptx::tmem_dealloc(tmem, 512);
  }
}

}  // namespace thin
}  // namespace vex
