// conv3x3 (pad 1, stride 1) of the Real-ESRGAN RRDB network (reference models/enhancement/vsr/realesrgan_handler.py:23-100)
// as an implicit GEMM on the sm_100a tensor cores: tcgen05.mma (fp16 x fp16 -> fp32 in TMEM), TMA-staged activation halo
// tiles, mbarrier rings.  One kernel serves every conv3x3 of the network except the 3-channel ends, 32 output channels
// per launch (the 64-channel layers are two launches):
//
//   ResidualDenseBlock.conv1..conv4 (:41-44, 50-53)  Cin = 64/96/128/160 -> 32, LeakyReLU(0.2), written into the channel
//        SLICE [64+32k, 96+32k) of the block's 192-channel NHWC "dense" buffer, so torch.cat((x, x1, ...), 1) costs nothing;
//   ResidualDenseBlock.conv5 (:45, 54-56)            192 -> 64, `x5 * 0.2 + x`, and for the third block of an RRDBBlock also
//        `out * 0.2 + x` of :36, both residuals applied in fp32 in the epilogue;
//   conv_body (:65, 88-89)                           64 -> 64, `feat + body_feat`;
//   conv_up1 / conv_up2 / conv_hr (:70-77, 93-98)    64 -> 64, LeakyReLU(0.2).
//   F.interpolate(scale_factor=2, mode='nearest') in front of conv_up1 / conv_up2 (:93-96) is done by the PRODUCER of the
//   tensor: its epilogue stores every output pixel to the 2 x 2 block of the upsampled buffer (`ups2`).
//
// Precision: fp16 operands and activations (the 23-block residual trunk needs the 11-bit mantissa: emulating the rounding
// points on the CPU gives 0.5 LSB of max error with fp16 against 4.8 LSB with bf16), fp32 accumulation, fp32
// epilogue arithmetic, saturating fp16 stores (a stale or overflowing value can never become Inf/NaN in a buffer).
//
// GEMM view.  One TMA box load brings a halo tile [ROWS+2][PITCH = 32 px][64 ch] (128 B per pixel, SWIZZLE_128B,
// out-of-bounds = 0 = conv padding).  An accumulator covers 4 tile rows x 32 tile columns = 128 GEMM rows, one warp-wide
// TMEM lane quarter per tile row.  The first version ran one MMA of N = 32 per tap (the A descriptor advanced by
// dy*32+dx rows): an N = 32 MMA reads 4 KB of A and 1 KB of B from shared memory and is bound by that read, 44.7 clocks
// against 16 of math (tools/ubench/mma_rate.cu), so the tensor pipe could never be more than a third busy.  Now the three
// HORIZONTAL taps sit side by side in N:
//        D[pixel m][dx*32 + co] = sum over dy, ci of  X[m + 32*dy][ci] * W[co][ci][dy][dx]
// - only the vertical tap moves the A descriptor (by whole tile rows, which also keeps it on the 1 KB swizzle atom) - and
// the epilogue finishes the convolution with two lane shuffles: the 32 lanes of a warp ARE the 32 columns of one tile row,
// so   out[x] = D[x-1][0:32] + D[x][32:64] + D[x+1][64:96].   A K = 16 step of an accumulator is 3 MMAs of N = 96 (56 clocks
// each, measured) instead of 9 of N = 32 (45 each): 168 clocks instead of 402.  30 of the 32 columns are valid outputs.
// Cin is consumed in 64-channel chunks; Cin = 96 / 160 issue only the two K = 16 steps of their last chunk that exist.
//
// Roles (384 threads, one CTA per SM, persistent over work items = (image, 8-row x 30-column tile)): warp 0 activation
// (TMA) producer; warps 1 and 7 MMA issuers, one accumulator each - a single thread needs ~65 clocks of its own
// instruction stream per tcgen05.mma here, more than the pipe needs for the MMA (measured with the MMAs switched off,
// tools/rconv_probe.py); warps 2-5 and 8-11 epilogue, one accumulator per group of four (TMEM lane quarter = warp & 3);
// warp 6 loads the layer's weights, which always
// fit (<= 3 chunks x 3 vertical taps x 12 KB) and stay resident: once per CTA, waited for during the first item only (even
// an already-complete mbarrier wait costs ~100 clocks of the issuing thread).  Two TMEM slots of two accumulators: the
// epilogue of item k overlaps the MMAs of item k+1.
#pragma once

#include <cuda_fp16.h>

#include "ptx.cuh"

namespace vex {
namespace rrdb {

constexpr int PITCH = 32;
constexpr int TILE_W = 30;
constexpr int KC = 64;
constexpr int SA_MAX = 4;        // activation stages: as many as fit beside the layer's weights (Params::sa), 2..4
constexpr int NTHREADS = 384;
constexpr int APS = 2;           // accumulators (128 pixels each) per work item
constexpr int ROWS = APS * 4;
constexpr int CO = 32;           // output channels per launch
constexpr int NB = 3 * CO;       // MMA N: the three horizontal taps of CO channels side by side
constexpr int ACC_COLS = 128;    // TMEM columns reserved per accumulator (NB used)
constexpr int MAXB = 16;         // images of different valid sizes per launch (canvas batching)
constexpr int MAX_NC = 3;        // 64-channel chunks of the widest input (192)

// work items of a vh x vw image on an H x W canvas: the valid part and, where the canvas has room, one more row / column
// (stored as zeros: the conv padding of the next layer)
__host__ __device__ constexpr int items_x(int vw, int W) { return ((vw + 1 < W ? vw + 1 : W) + TILE_W - 1) / TILE_W; }
__host__ __device__ constexpr int items_y(int vh, int H) { return ((vh + 1 < H ? vh + 1 : H) + ROWS - 1) / ROWS; }

struct Cfg {
  static constexpr int A_STAGE = (ROWS + 2) * PITCH * 128;
  static constexpr int W_SLICE = NB * 128;              // one (chunk, vertical tap) slice: 96 rows (dx*32 + co) x 64 fp16
  static constexpr int SW = MAX_NC * 3;                 // every slice of the widest layer
  static constexpr int BAR_BYTES = 512;                 // a_full/a_empty[SA_MAX], w_full[SW], acc_full/acc_empty[2], TMEM slot
  static constexpr int TAIL_BYTES = BAR_BYTES + 256 + 8 * 2048;   // barriers, bias, the 8 epilogue warps' transpose stages
  static constexpr int SMEM_MAX = 227 * 1024;
  static constexpr int TMEM_COLS = 2 * APS * ACC_COLS;  // 512
  // shared memory of a launch: [sa activation stages][NC*3 weight slices][barriers][bias][stages], + 1 KB for the 1024-byte
  // alignment of the base.  More activation stages = more bytes in flight per SM: a layer's input comes from HBM (796 MB
  // per dense buffer at 1080p) and since the MMAs got 2.4x shorter the loads are what the wide layers wait for.
  __host__ __device__ static constexpr int stages(int nslice) {
    const int room = (SMEM_MAX - 1024 - TAIL_BYTES - nslice * W_SLICE) / A_STAGE;
    return room > SA_MAX ? SA_MAX : room;
  }
  __host__ __device__ static constexpr int smem(int sa, int nslice) { return 1024 + sa * A_STAGE + nslice * W_SLICE + TAIL_BYTES; }
};
static_assert(Cfg::stages(Cfg::SW) >= 2, "shared memory");

struct Params {
  const uint8_t* wpack;   // [chunk][dy] slices of NB x 128 B (row dx*32 + co), SWIZZLE_128B image, fp16: this launch's 32 output channels
  const float* bias;      // [CO]
  __half* out;            // NHWC [n][H][W][out_cs] (ups2: [n][2H][2W][out_cs]); channels [out_c0, out_c0 + CO) are written
  const __half* r1;       // optional residual, NHWC with channel stride r1_cs, channels [0, CO):  v = v * a1 + r1
  const __half* r2;       // optional second residual:                                             v = v * a2 + r2
  float a1, a2;
  int out_cs, out_c0, r1_cs, r2_cs;
  int lrelu;              // LeakyReLU(0.2) after the bias
  int ups2;               // store every pixel to the 2 x 2 block of a twice as large output (nearest-neighbour upsample)
  int sa;                 // activation stages of this launch (Cfg::stages(NC * 3), or fewer: option rrdb_stages)
  int H, W, NC;           // NC = 64-channel chunks of the input that are read
  int klast;              // K = 16 steps of the LAST chunk: 2 when Cin % 64 == 32 (conv2 / conv4 of a dense block), else 4
  int half_last;          // klast == 2 and the last chunk is loaded through `hmap`: 32 channels per pixel (64-byte rows, SWIZZLE_64B),
                          // half the bytes of a full chunk - these layers wait for their activations, and the other 32 channels
                          // of that chunk are the slice this very layer is writing
  int tiles_x, tiles_y, n_img;
  int nblk;               // 1, or 2: both 32-channel halves of a 64-channel layer in ONE launch.  Work item it = 2 * tile + half and the
                          // grid is even, so even CTAs only ever see half 0 and odd CTAs half 1 (each keeps its own half's weights
                          // resident) while neighbouring CTAs work on the same tile at the same time: the second read of its
                          // activations hits L2 instead of HBM.  wpack / bias / out_c0 / r1 / r2 advance by one half per blk.
  int reverse;            // sweep the work items from the last to the first: consecutive layers alternate, so a layer starts on the
                          // ~100 MB its predecessor touched last, which are still in L2 (a dense buffer is 0.8 GB at 1080p)
  // Canvas batching (the tile loop of `_tile_process`, realesrgan_handler.py:324-372, in ONE launch per layer): the n_tab
  // images of the batch share the H x W canvas but image i is only vh[i] x vw[i] large, anchored at (0, 0).  Work items
  // cover the valid part plus one more row / column, and every store outside the valid part writes ZERO, so the next
  // layer sees exactly the zero padding a tile run alone gets from the tensor map's out-of-bounds fill.  n_tab == 0: all
  // n_img images are H x W (tiles_x / tiles_y above).
  int n_tab;
  short vh[MAXB], vw[MAXB];
  int item0[MAXB + 1];    // first work item of image i; item0[n_tab] = total
  int probe;              // -DVEX_DEBUG_PROBES builds only (wrong pixels, timing experiments): 256 = the activation producer loads the
                          // first SA stages and then only signals; 512 = the epilogue only waits and releases; 2048 = no MMAs
  long long* dbg;         // optional [gridDim.x][8] cycle counters (option conv_dbg): MMA thread 0 total, 1 wait acc_empty, 2 wait a_full,
                          // 3 wait w_full; epilogue warp 2: 4 wait acc_full, 5 work; producer: 6 wait a_empty
};

// Programmatic dependent launch: the ~350 layer kernels of one pass form a chain on one stream.  A kernel launched with
// the programmatic-serialisation attribute may start while its predecessor's last CTAs are still running; its prologue
// (barrier init, TMEM allocation, bias, tensor-map prefetch, weights) then overlaps the predecessor's tail, and `pdl_wait`
// holds every access to activations until the predecessor has completed and flushed.  `pdl_launch_dependents` lets the
// successor start as early as possible.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// fp32 pair -> packed fp16, round to nearest, saturating to +-65504
__device__ __forceinline__ uint32_t pack_f16x2_sat(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float2 unpack_f16x2(uint32_t w) {
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}
struct Item { int n, x0, y0, vh, vw; };
// work item -> image, tile origin and the image's valid size
template <class P>
__device__ __forceinline__ Item decode_item(const P& p, int total, int it) {
  Item r;
  if (p.nblk == 2) { it >>= 1; total >>= 1; }           // the tile; the half is the CTA's parity
  if (p.reverse) it = total - 1 - it;
  if (p.n_tab == 0) {
    const int q = it / p.tiles_x;
    r.n = q / p.tiles_y; r.x0 = (it % p.tiles_x) * TILE_W; r.y0 = (q % p.tiles_y) * ROWS; r.vh = p.H; r.vw = p.W;
  } else {
    int n = 0;
    while (n + 1 < p.n_tab && it >= p.item0[n + 1]) ++n;
    r.vh = p.vh[n]; r.vw = p.vw[n];
    const int tx = items_x(r.vw, p.W), l = it - p.item0[n];
    r.n = n; r.x0 = (l % tx) * TILE_W; r.y0 = (l / tx) * ROWS;
  }
  return r;
}

// instruction descriptor for kind::f16 with fp16 A/B (format 0), fp32 D, both K-major
__host__ __device__ constexpr uint32_t idesc_f16_f32(int m, int n) {
  return (1u << 4) | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

__global__ void __launch_bounds__(NTHREADS, 1)
rconv3x3_kernel(const __grid_constant__ CUtensorMap amap, const __grid_constant__ CUtensorMap hmap,
                const __grid_constant__ Params p) {
  using C = Cfg;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);

  const uint32_t SA = (uint32_t)p.sa;
  const uint32_t a_stage0 = sbase;
  const uint32_t w_stage0 = sbase + SA * C::A_STAGE;
  const uint32_t bar_off = SA * C::A_STAGE + (uint32_t)p.NC * 3 * C::W_SLICE;
  const uint32_t bar0 = sbase + bar_off;
  // barrier map (8 B each): a_full[SA_MAX] a_empty[SA_MAX] w_full[SW] acc_full[2] acc_empty[2] | tmem slot
  const uint32_t a_full = bar0, a_empty = bar0 + 8 * SA_MAX;
  const uint32_t w_full = a_empty + 8 * SA_MAX;
  const uint32_t acc_full = w_full + 8 * C::SW, acc_empty = acc_full + 16;
  const uint32_t tmem_slot = acc_empty + 16;
  static_assert(16 * SA_MAX + 8 * C::SW + 32 + 8 <= C::BAR_BYTES, "barrier block");
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + bar_off + (tmem_slot - bar0));
  float* bias_s = reinterpret_cast<float*>(sm + bar_off + C::BAR_BYTES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = (p.n_tab ? p.item0[p.n_tab] : p.tiles_x * p.tiles_y * p.n_img) * p.nblk;
  const int blk = p.nblk == 2 ? (int)(blockIdx.x & 1u) : 0;

  if (threadIdx.x == 0) {
    // consumer-side barriers count APS arrivals: each of the two MMA threads commits its own MMAs
    for (int i = 0; i < SA_MAX; ++i) { ptx::mbar_init(a_full + 8 * i, 1); ptx::mbar_init(a_empty + 8 * i, APS); }
    for (int i = 0; i < C::SW; ++i) ptx::mbar_init(w_full + 8 * i, 1);
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(acc_full + 8 * i, APS); ptx::mbar_init(acc_empty + 8 * i, 8); }
    ptx::fence_mbar_init();
    ptx::prefetch_tensormap(&amap);
    if (p.half_last) ptx::prefetch_tensormap(&hmap);
  }
  if (threadIdx.x < CO) bias_s[threadIdx.x] = p.bias[blk * CO + threadIdx.x];
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;
  pdl_launch_dependents();                              // the next layer may begin its prologue

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (ptx::elect_one()) {
      uint32_t sa = 0, pa = 0;
      const bool dbg = p.dbg != nullptr;
      long long t6 = 0, tq = 0;
#ifdef VEX_DEBUG_PROBES
      int nload = 0;
#endif
      pdl_wait();                                       // activations of the previous layer
      for (int it = blockIdx.x; it < total; it += gridDim.x) {
        const Item im = decode_item(p, total, it);
        const int x0 = im.x0, y0 = im.y0, n = im.n;
        for (int c = 0; c < p.NC; ++c) {
          if (dbg) tq = clock64();
          ptx::mbar_wait(a_empty + 8 * sa, pa ^ 1);
          if (dbg) t6 += clock64() - tq;
#ifdef VEX_DEBUG_PROBES
          if ((p.probe & 256) && nload >= SA) { ptx::mbar_arrive(a_full + 8 * sa); if (++sa == SA) { sa = 0; pa ^= 1; } continue; }
          ++nload;
#endif
          if (p.half_last && c == p.NC - 1) {
            ptx::mbar_expect_tx(a_full + 8 * sa, C::A_STAGE / 2);
            ptx::tma_load_4d(a_stage0 + sa * C::A_STAGE, &hmap, a_full + 8 * sa, c * KC, x0 - 1, y0 - 1, n);
          } else {
            ptx::mbar_expect_tx(a_full + 8 * sa, C::A_STAGE);
            ptx::tma_load_4d(a_stage0 + sa * C::A_STAGE, &amap, a_full + 8 * sa, c * KC, x0 - 1, y0 - 1, n);
          }
          if (++sa == SA) { sa = 0; pa ^= 1; }
        }
      }
      if (dbg) p.dbg[blockIdx.x * 8 + 6] = t6;
    }
  } else if (warp == 6) {
    // ============================== weight producer ===========================
    // Weights do not depend on the previous layer: no pdl_wait here, the slices land while the predecessor drains.
    if (ptx::elect_one()) {
      const int nslice = p.NC * 3;
      for (int i = 0; i < nslice; ++i) {
        ptx::mbar_expect_tx(w_full + 8 * i, C::W_SLICE);
        ptx::bulk_load(w_stage0 + i * C::W_SLICE, p.wpack + (size_t)(blk * nslice + i) * C::W_SLICE, C::W_SLICE, w_full + 8 * i);
      }
    }
  } else if (warp == 1 || warp == 7) {
    // ============================== MMA issuers ===============================
    static_assert(APS == 2, "one MMA warp per accumulator");
    const int a = warp == 1 ? 0 : 1;
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = idesc_f16_f32(128, NB);
      uint32_t sa = 0, pa = 0, kk = 0;
      const bool dbg = p.dbg != nullptr && a == 0;
      long long t1 = 0, t2 = 0, t3 = 0, tq = 0;
      const long long tm0 = dbg ? clock64() : 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x, ++kk) {
        const uint32_t slot = kk & 1, pacc = (kk >> 1) & 1;
        const uint32_t dcol = tmem + (slot * APS + a) * ACC_COLS;
        if (dbg) tq = clock64();
        ptx::mbar_wait(acc_empty + 8 * slot, pacc ^ 1);   // the epilogue has drained this slot (first use passes at once)
        if (dbg) t1 += clock64() - tq;
        ptx::tc_fence_after();
        for (int c = 0; c < p.NC; ++c) {
          if (dbg) tq = clock64();
          ptx::mbar_wait(a_full + 8 * sa, pa);
          if (dbg) t2 += clock64() - tq;
          // accumulator a = tile rows 4a .. 4a+3; vertical tap dy reads tile rows 4a+dy .. 4a+dy+3 (the box starts at y0 - 1)
          const bool half = p.half_last && c == p.NC - 1;     // 64-byte rows
          const uint32_t rowb = half ? 64u : 128u;
          const uint32_t abase = a_stage0 + sa * C::A_STAGE + (4 * a) * PITCH * rowb;
          const int ksteps = c == p.NC - 1 ? p.klast : KC / 16;
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            const uint32_t sl = c * 3 + dy;
            if (kk == 0) {                                // resident weights: phase 0 of the slice's barrier stays complete
              if (dbg) tq = clock64();
              ptx::mbar_wait(w_full + 8 * sl, 0);
              if (dbg) t3 += clock64() - tq;
            }
            ptx::tc_fence_after();
            const uint64_t ad0 = half ? ptx::smem_desc_sw64(abase + dy * PITCH * 64) : ptx::smem_desc_sw128(abase + dy * PITCH * 128, 0);
            const uint64_t bd0 = ptx::smem_desc_sw128(w_stage0 + sl * C::W_SLICE, 0);
#pragma unroll
            for (int k = 0; k < KC / 16; ++k) {
              if (k >= ksteps) break;
#ifdef VEX_DEBUG_PROBES
              if (p.probe & 2048) continue;               // no MMAs at all: what the other roles cost on their own
#endif
              ptx::mma_bf16_ss(dcol, ptx::desc_advance(ad0, k * 32), ptx::desc_advance(bd0, k * 32), idesc, (c | dy | k) != 0);
            }
          }
          ptx::tc_commit(a_empty + 8 * sa);
          if (++sa == SA) { sa = 0; pa ^= 1; }
        }
        ptx::tc_commit(acc_full + 8 * slot);
      }
      if (dbg) { long long* d = p.dbg + blockIdx.x * 8; d[0] = clock64() - tm0; d[1] = t1; d[2] = t2; d[3] = t3; }
    }
  } else {
    // ============================== epilogue ==================================
    // 8 warps: warps 2-5 drain accumulator 0, warps 8-11 accumulator 1 (since the MMAs got 2.4x shorter, four warps were
    // what the 64-channel-input layers and the heads waited for).  TMEM lane quarter q = warp & 3 = tile row 4a + q of
    // accumulator a; lane = tile column (pixel x0 - 1 + lane).
    // A lane reads its three 32-column groups and finishes the horizontal taps with its neighbours:
    //   out(lane) = D[lane-1][0:32] + D[lane][32:64] + D[lane+1][64:96]          (lanes 1..30 are outputs)
    // then bias, LeakyReLU, residuals (fp32) -> 64 bytes of fp16.  Global accesses go through a 2 KB per-warp stage so
    // that one instruction covers 8 pixels x 64 B (lane l <-> pixel 8i + (l >> 2), 16-byte chunk l & 3).  Stage layout:
    // pixel px at px*64, chunk j at (j ^ ((px >> 1) & 3)) * 16.
    const int q = warp & 3, a = warp >= 8 ? 1 : 0;
    uint8_t* stage = sm + bar_off + C::BAR_BYTES + 256 + (warp >= 8 ? warp - 4 : warp - 2) * 2048;
    const uint32_t tq = tmem + ((uint32_t)(q * 32) << 16);
    const int cpx = lane >> 2, cch = lane & 3;
    int own_off[4], co_off[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      own_off[j] = lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4);
      const int px = 8 * j + cpx;
      co_off[j] = px * 64 + ((cch ^ ((px >> 1) & 3)) << 4);
    }
    const int oh = p.ups2 ? 2 * p.H : p.H, ow = p.ups2 ? 2 * p.W : p.W;
    const int oc0 = p.out_c0 + blk * CO;
    const __half* r1p = p.r1 ? p.r1 + blk * CO : nullptr;
    const __half* r2p = p.r2 ? p.r2 + blk * CO : nullptr;
    uint32_t kk = 0;
    pdl_wait();                                         // residuals are read and the output buffer is written below
    for (int it = blockIdx.x; it < total; it += gridDim.x, ++kk) {
      const uint32_t slot = kk & 1, pacc = (kk >> 1) & 1;
      const Item im = decode_item(p, total, it);
      const int xl = im.x0 - 1, n = im.n;               // x of tile column 0
      const int ybase = im.y0 + q;                      // accumulator row m = q*32 + lane of accumulator a -> y = ybase + 4a
      bool pxok[4], pxval[4];                           // an output inside the canvas (stored) / inside the image (else stored as zero)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int px = 8 * i + cpx;
        pxok[i] = px >= 1 && px <= TILE_W && xl + px < p.W;
        pxval[i] = xl + px < im.vw;
      }
      const bool dbg = p.dbg != nullptr && warp == 2 && lane == 0;
      const long long te0 = dbg ? clock64() : 0;
      ptx::mbar_wait(acc_full + 8 * slot, pacc);
      const long long te1 = dbg ? clock64() : 0;
      ptx::tc_fence_after();
#ifdef VEX_DEBUG_PROBES
      const int nunits = (p.probe & 512) ? 0 : 1;
#else
      constexpr int nunits = 1;
#endif
#pragma unroll 1
      for (int u = 0; u < nunits; ++u) {
        const int y = ybase + 4 * a;
        const bool yok = y < p.H, yval = y < im.vh;
        uint32_t v0[32], v1[32], v2[32];
        const uint32_t dcol = tq + (slot * APS + a) * ACC_COLS;
        ptx::tmem_ld_32x32(dcol, v0);
        ptx::tmem_ld_32x32(dcol + CO, v1);
        ptx::tmem_ld_32x32(dcol + 2 * CO, v2);
        // residual rows in the co-pattern (pixel 8i + cpx, chunk cch), issued before the TMEM wait
        uint4 s1[4], s2[4];
        const long long pix0 = ((long long)n * p.H + y) * p.W + xl + cpx;
        if (p.r1) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            s1[i] = (yval && pxok[i] && pxval[i]) ? __ldg(reinterpret_cast<const uint4*>(r1p + (pix0 + 8 * i) * p.r1_cs + cch * 8))
                                                  : make_uint4(0, 0, 0, 0);
        }
        if (p.r2) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            s2[i] = (yval && pxok[i] && pxval[i]) ? __ldg(reinterpret_cast<const uint4*>(r2p + (pix0 + 8 * i) * p.r2_cs + cch * 8))
                                                  : make_uint4(0, 0, 0, 0);
        }
        ptx::tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
#ifdef VEX_DEBUG_PROBES
          const bool noshfl = p.probe & 4096;
          const float left = noshfl ? __uint_as_float(v0[j]) : __shfl_up_sync(0xffffffffu, __uint_as_float(v0[j]), 1);
          const float right = noshfl ? __uint_as_float(v2[j]) : __shfl_down_sync(0xffffffffu, __uint_as_float(v2[j]), 1);
#else
          const float left = __shfl_up_sync(0xffffffffu, __uint_as_float(v0[j]), 1);      // D[lane-1][j]       (tap dx = 0)
          const float right = __shfl_down_sync(0xffffffffu, __uint_as_float(v2[j]), 1);   // D[lane+1][64 + j]  (tap dx = 2)
#endif
          float t = (left + __uint_as_float(v1[j])) + right + bias_s[j];
          if (p.lrelu) t = t > 0.f ? t : 0.2f * t;        // nn.LeakyReLU(negative_slope=0.2)
          f[j] = t;
        }
        if (p.r1) {                                       // x5 * 0.2 + x  (:57)  /  feat + body_feat  (:89, a1 = 1)
#pragma unroll
          for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(stage + co_off[i]) = s1[i];
          __syncwarp();
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            const uint4 w = *reinterpret_cast<const uint4*>(stage + own_off[j4]);
            const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 rr = unpack_f16x2(ww[e]);
              f[j4 * 8 + 2 * e] = fmaf(f[j4 * 8 + 2 * e], p.a1, rr.x);
              f[j4 * 8 + 2 * e + 1] = fmaf(f[j4 * 8 + 2 * e + 1], p.a1, rr.y);
            }
          }
          __syncwarp();
        }
        if (p.r2) {                                       // out * 0.2 + x of the enclosing RRDBBlock (:36)
#pragma unroll
          for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(stage + co_off[i]) = s2[i];
          __syncwarp();
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            const uint4 w = *reinterpret_cast<const uint4*>(stage + own_off[j4]);
            const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 rr = unpack_f16x2(ww[e]);
              f[j4 * 8 + 2 * e] = fmaf(f[j4 * 8 + 2 * e], p.a2, rr.x);
              f[j4 * 8 + 2 * e + 1] = fmaf(f[j4 * 8 + 2 * e + 1], p.a2, rr.y);
            }
          }
          __syncwarp();
        }
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4)
          *reinterpret_cast<uint4*>(stage + own_off[j4]) =
              make_uint4(pack_f16x2_sat(f[j4 * 8 + 0], f[j4 * 8 + 1]), pack_f16x2_sat(f[j4 * 8 + 2], f[j4 * 8 + 3]),
                         pack_f16x2_sat(f[j4 * 8 + 4], f[j4 * 8 + 5]), pack_f16x2_sat(f[j4 * 8 + 6], f[j4 * 8 + 7]));
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 o = *reinterpret_cast<const uint4*>(stage + co_off[i]);
#ifdef VEX_DEBUG_PROBES
          if (p.probe & 8192) continue;                   // no global stores
#endif
          if (!(yok && pxok[i])) continue;
          if (!(yval && pxval[i])) o = make_uint4(0, 0, 0, 0);
          const int x = xl + 8 * i + cpx;
          if (!p.ups2) {
            *reinterpret_cast<uint4*>(p.out + (((size_t)n * oh + y) * ow + x) * p.out_cs + oc0 + cch * 8) = o;
          } else {
#pragma unroll
            for (int d = 0; d < 4; ++d)
              *reinterpret_cast<uint4*>(p.out + (((size_t)n * oh + 2 * y + (d >> 1)) * ow + 2 * x + (d & 1)) * p.out_cs +
                                        oc0 + cch * 8) = o;
          }
        }
        __syncwarp();
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc_empty + 8 * slot);
      if (dbg) { p.dbg[blockIdx.x * 8 + 4] += te1 - te0; p.dbg[blockIdx.x * 8 + 5] += clock64() - te1; }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    ptx::tmem_dealloc(tmem, C::TMEM_COLS);
  }
}

// nn.Conv2d(3, 64) of conv_first and nn.Conv2d(64, 3) of conv_last run on the thin-layer kernels of thin_tc.cuh
// (im2col27_kernel<MODE_FIRST64, 64>, gemm27_kernel<MODE_LAST_F16>).

}  // namespace rrdb
}  // namespace vex
