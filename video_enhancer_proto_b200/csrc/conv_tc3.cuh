// conv3x3 (pad 1, stride 1) + folded-BN bias + ReLU (+ post-ReLU skip) as an implicit GEMM on the sm_100a
// tensor cores: tcgen05.mma (bf16 x bf16 -> fp32 in TMEM), TMA-staged activation halo tiles, mbarrier rings.
// Replaces conv2/bn2/relu(+x1), conv3/bn3/relu, conv4/bn4/relu of AdvancedUpscaler (reference app.py:292-294).
//
// GEMM view: M = output pixels, N = Cout block (BN), K = 9 taps x Cin.  Activations are NHWC bf16.
//
// The trick that avoids loading the activations 9 times (once per tap): one TMA box load brings a halo tile
//   [ROWS+2 rows][PITCH=32 pixels][64 channels]  (128 B per pixel, 128B-swizzled, OOB = 0 = conv padding)
// into shared memory.  Pixels are then simply consecutive 128-byte rows of a K-major SWIZZLE_128B operand, so
// the A operand of tap (dy,dx) is THE SAME buffer with its start address advanced by (dy*32+dx) rows.  The
// M index runs over the flat [row][32] grid; the last two columns of each row are computed on wrapped data
// and discarded (30 valid output columns per 32: 6.25 % redundant MMA work, zero redundant loads).
//
// CTA = 10 warps: warp 0 TMA producer (1 lane), warp 1 MMA issuer (1 lane) + TMEM owner, warps 2-9 epilogue
// (TMEM lane quarter = warp & 3; two warps per quarter split the 32-column groups).
//
// v2 (this file): the 512 TMEM columns are TWO accumulator slots (BN=128: 2 x 2 accumulators = 256 pixels per
// item, BN=256: 2 x 1 = 128 pixels), so the epilogue of item k drains slot k&1 while the MMAs of item k+1 fill
// the other slot (v1 used all columns for one item and measured 13-34 % of the MMA thread's time waiting for the
// epilogue, profiles/).  Halving the pixels per item would double the weight traffic per pixel (every weight
// slice is re-streamed from L2 for every item), so CTAs run as CLUSTERS OF TWO that consume the same slice
// sequence in lock-step: each CTA fetches half of every slice and multicasts it to both
// (cp.async.bulk ... .multicast::cluster), and a ring slot is recycled when the MMAs of BOTH CTAs have retired
// (tcgen05.commit ... .multicast::cluster onto a 2-arrival mbarrier).  Per-SM L2 weight traffic stays what v1 had.
// Persistent grid; work item = (image, 8- or 4-row x 30-column tile); both CTAs of a pair run the same number
// of items (the last one may be a repeat with its stores masked).
//
// v3 (this file): the pair issues ONE tcgen05.mma.cta_group::2 (M = 256 = the two CTAs' pixel tiles, N = BN) from
// the leader CTA.  Each CTA keeps only its HALF of every weight slice (BN/2 rows) in shared memory and the tensor
// cores exchange the halves, so an SM reads 4 KB (A) + 2 KB (B/2) per K=16 step instead of 8 KB: the SS MMA of v1/v2
// saturates the 128 B/clk shared-memory port at N = 128 (measured: profiles/), v3 does not.  No multicast is needed.
// Conventions measured with tools/ubench/mma2sm.cu: CTA r holds A rows [128r, 128r+128) and B rows [N/2 r, N/2 r + N/2)
// at the same shared offsets, D lands in each CTA's own 128 TMEM lanes.  The peer's otherwise idle MMA warp relays
// "my stage is full" to the leader with remote mbarrier arrives; the leader's tcgen05.commit.cta_group::2 multicast
// frees stages and publishes accumulators in both CTAs; the peer's epilogue warps hand the slot back remotely.
#pragma once

#include "ptx.cuh"

namespace vex {
namespace convtc3 {

constexpr int PITCH = 32;     // pixels per smem tile row (30 valid outputs + 2 halo)
constexpr int TILE_W = 30;
constexpr int KC = 64;        // channels per K chunk = one 128 B swizzle row
constexpr int SA = 2;         // activation stages
constexpr int NTHREADS = 352;    // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue, warp 10 second MMA issuer (BN = 128 only)
constexpr int ISS2_WARP = 10;
// Two issuing threads when an item has two accumulators (BN = 128: conv2, conv4), one accumulator each.  A single thread
// spends ~55-65 clocks of its own instruction stream per tcgen05.mma (measured in rrdb_tc.cuh with the MMAs switched off) plus
// two barrier waits per tap; an N = 128 MMA of a CTA pair takes the pipe 64 clocks, so one issuer was right at the limit
// (tensor pipe 87.6 % active in conv2, 97 % in conv4) while the N = 256 kernel (128 clocks per MMA) reached 99.9 %.
// Both issuers wait on the same "full" barriers; the barriers they commit to count one arrival per issuer.
// This loop's code generation decides the speed of the BN = 128 layers: the same split with run-time accumulator bounds,
// with the loop instantiated per accumulator in a generic lambda, or with the tap loop unrolled each measured conv4 at
// 7.8-9 ms instead of 6.1 (tools/conv_iss_ab.py) - change it only with that tool at hand.

template <int BN>
struct Cfg {
  static_assert(BN == 128 || BN == 256, "BN");
  static constexpr int NACC = 512 / BN;                 // accumulators in TMEM
  static constexpr int APS = NACC / 2;                  // accumulators per slot (= per work item)
  static constexpr int ROWS = APS * 4;                  // output rows per work item (128 px = 4 rows of 32)
  static constexpr int A_STAGE = (ROWS + 2) * PITCH * 128;
  static constexpr int W_SLICE = BN * 128;              // one (chunk, tap) weight slice in global: BN rows x 64 bf16
  static constexpr int W_STAGE = W_SLICE / 2;           // this CTA's half of it (BN/2 rows)
  static constexpr int SW = BN == 128 ? 14 : 8;
  static constexpr int BAR_OFF = SA * A_STAGE + SW * W_STAGE;
  static constexpr int BIAS_OFF = BAR_OFF + 512;
  static constexpr int STAGE_OFF = BIAS_OFF + BN * 4;   // 8 epilogue warps x 2 KB transpose stage
  static constexpr int SMEM = STAGE_OFF + 8 * 2048 + 1024; // + alignment slack
};

struct Params {
  const uint8_t* wpack;   // [cout_block][chunk][tap] slices, each BN x 128 B, pre-swizzled (SWIZZLE_128B image)
  const float* bias;      // [Cout] folded conv bias + BN shift
  __nv_bfloat16* out;     // NHWC [n][H][W][Cout]
  const __nv_bfloat16* skip;  // NHWC [n][H][W][Cout] added AFTER the ReLU (app.py:292), or nullptr
  int H, W, NC, Cout;     // NC = Cin / 64
  int tiles_x, tiles_y, n_img;
  int desc_base_offset;   // debug knob, see ptx::smem_desc_sw128
  int epi_variant;        // bring-up timing probes (WRONG RESULTS): 1 = no global stores, 2 = no tcgen05.ld, 4 = no skip loads
  long long* dbg;         // optional [gridDim.x][8] cycle counters (MMA thread: total, wait_acc, wait_a, wait_w; epilogue: wait_full, work)
};

// DIRECT: the epilogue reads the skip and writes the output straight from/to global (64 contiguous bytes per
// lane) instead of transposing through shared memory.  The SS MMAs of this kernel already consume the whole
// shared-memory bandwidth (M=128, N=128: 8 KB of operands per 64 tensor cycles), so staging traffic costs MMA time.
template <int BN, bool SKIP, bool DIRECT>
__global__ void __launch_bounds__(NTHREADS, 1)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap amap, const Params p) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;        // SWIZZLE_128B atoms need 1024 B alignment
  uint8_t* sm = smem_raw + (sbase - raw);

  const uint32_t a_stage0 = sbase;
  const uint32_t w_stage0 = sbase + SA * C::A_STAGE;
  const uint32_t bar0 = sbase + C::BAR_OFF;
  // barrier map (8 B each): a_full[SA] a_empty[SA] w_full[SW] w_empty[SW] acc_full[2] acc_empty[2] | tmem slot
  const uint32_t a_full = bar0, a_empty = bar0 + 8 * SA;
  const uint32_t w_full = a_empty + 8 * SA, w_empty = w_full + 8 * C::SW;
  const uint32_t acc_full = w_empty + 8 * C::SW, acc_empty = acc_full + 16;
  const uint32_t a_peer = acc_empty + 16, w_peer = a_peer + 8 * SA;     // leader only: the peer's stage is full
  const uint32_t tmem_slot = w_peer + 8 * C::SW;
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + C::BAR_OFF + (tmem_slot - bar0));
  float* bias_s = reinterpret_cast<float*>(sm + C::BIAS_OFF);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nb = blockIdx.y;                            // Cout block
  const int total = p.tiles_x * p.tiles_y * p.n_img;
  const int iters = (total + (int)gridDim.x - 1) / (int)gridDim.x;   // identical in both CTAs of a pair
  const uint32_t rank = ptx::cluster_ctarank();

  if (threadIdx.x == 0) {
    constexpr int NISS = C::APS == 2 ? 2 : 1;           // issuing threads (in the leader CTA; their commits are multicast to both)
    for (int i = 0; i < SA; ++i) { ptx::mbar_init(a_full + 8 * i, 1); ptx::mbar_init(a_empty + 8 * i, NISS); }
    for (int i = 0; i < C::SW; ++i) { ptx::mbar_init(w_full + 8 * i, 1); ptx::mbar_init(w_empty + 8 * i, NISS); ptx::mbar_init(w_peer + 8 * i, 1); }
    for (int i = 0; i < SA; ++i) ptx::mbar_init(a_peer + 8 * i, 1);
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(acc_full + 8 * i, NISS); ptx::mbar_init(acc_empty + 8 * i, 16); }   // 8 + 8 epilogue warps of the pair
    ptx::fence_mbar_init();
    ptx::prefetch_tensormap(&amap);
  }
  for (int i = threadIdx.x; i < BN; i += NTHREADS) bias_s[i] = p.bias[nb * BN + i];
  if (warp == 1) {
    ptx::tmem_alloc2(tmem_slot, 512);
    ptx::tmem_relinquish2();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();                                  // the peer's barriers exist before anything is multicast to them
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (ptx::elect_one()) {
      const uint8_t* wsrc = p.wpack + (size_t)nb * p.NC * 9 * C::W_SLICE + rank * C::W_STAGE;   // this CTA's rows of B
      uint32_t sa = 0, pa = 0, sw = 0, pw = 0;
      for (int kk = 0; kk < iters; ++kk) {
        const int it = min((int)blockIdx.x + kk * (int)gridDim.x, total - 1);
        const int txi = it % p.tiles_x, r = it / p.tiles_x;
        const int tyi = r % p.tiles_y, n = r / p.tiles_y;
        const int x0 = txi * TILE_W, y0 = tyi * C::ROWS;
        for (int c = 0; c < p.NC; ++c) {
          ptx::mbar_wait(a_empty + 8 * sa, pa ^ 1);
          ptx::mbar_expect_tx(a_full + 8 * sa, C::A_STAGE);
          ptx::tma_load_4d(a_stage0 + sa * C::A_STAGE, &amap, a_full + 8 * sa, c * KC, x0 - 1, y0 - 1, n);
          if (++sa == SA) { sa = 0; pa ^= 1; }
          for (int t = 0; t < 9; ++t) {
            ptx::mbar_wait(w_empty + 8 * sw, pw ^ 1);   // freed by the leader's commit (multicast to both CTAs)
            ptx::mbar_expect_tx(w_full + 8 * sw, C::W_STAGE);
            ptx::bulk_load(w_stage0 + sw * C::W_STAGE, wsrc + (size_t)(c * 9 + t) * C::W_SLICE, C::W_STAGE,
                           w_full + 8 * sw);
            if (++sw == C::SW) { sw = 0; pw ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1 || warp == ISS2_WARP) {
    // ============================== MMA issuer(s) =============================
    const bool second = warp == ISS2_WARP;              // takes accumulator 1 when the item has two
    const int a_lo = (C::APS == 2 && second) ? 1 : 0, a_hi = C::APS == 2 ? a_lo + 1 : C::APS;
    const bool elected = ptx::elect_one() && !(second && (C::APS != 2 || rank != 0));
    if (elected && rank != 0) {
      // ---- peer: relay "stage full" to the leader (same loop structure as the issuer below)
      uint32_t sa = 0, pa = 0, sw = 0, pw = 0;
      const uint32_t ra = ptx::mapa(a_peer, 0), rw = ptx::mapa(w_peer, 0);
      for (int kk = 0; kk < iters; ++kk) {
        for (int c = 0; c < p.NC; ++c) {
          ptx::mbar_wait(a_full + 8 * sa, pa);
          ptx::mbar_arrive_remote_relaxed(ra + 8 * sa);
          if (++sa == SA) { sa = 0; pa ^= 1; }
          for (int t = 0; t < 9; ++t) {
            ptx::mbar_wait(w_full + 8 * sw, pw);
            ptx::mbar_arrive_remote_relaxed(rw + 8 * sw);
            if (++sw == C::SW) { sw = 0; pw ^= 1; }
          }
        }
      }
    } else if (elected) {
      constexpr uint32_t idesc = ptx::idesc_bf16_f32(256, BN);
      uint32_t sa = 0, pa = 0, sw = 0, pw = 0;
      long long t_acc = 0, t_a = 0, t_w = 0, t0 = 0, t1;
      const bool dbg = p.dbg != nullptr && !second;
      const long long t_begin = dbg ? clock64() : 0;
      for (int kk = 0; kk < iters; ++kk) {
        const uint32_t slot = kk & 1, pacc = (kk >> 1) & 1;
        if (dbg) t0 = clock64();
        ptx::mbar_wait(acc_empty + 8 * slot, pacc);     // slot drained and re-initialised with the bias
        if (dbg) { t1 = clock64(); t_acc += t1 - t0; }
        ptx::tc_fence_after();
        for (int c = 0; c < p.NC; ++c) {
          if (dbg) t0 = clock64();
          ptx::mbar_wait(a_full + 8 * sa, pa);
          ptx::mbar_wait(a_peer + 8 * sa, pa);
          if (dbg) { t1 = clock64(); t_a += t1 - t0; }
          const uint32_t abase = a_stage0 + sa * C::A_STAGE;
          for (int t = 0; t < 9; ++t) {
            if (dbg) t0 = clock64();
            ptx::mbar_wait(w_full + 8 * sw, pw);
            if (dbg) { t1 = clock64(); t_a += t1 - t0; t0 = t1; }   // TEMP: t_a = local w wait, t_w = peer w wait
            ptx::mbar_wait(w_peer + 8 * sw, pw);
            if (dbg) { t1 = clock64(); t_w += t1 - t0; }
            ptx::tc_fence_after();
            const uint32_t wbase = w_stage0 + sw * C::W_STAGE;
            const uint32_t tap_off = ((t / 3) * PITCH + (t % 3)) * 128;
const uint64_t ad0 = ptx::smem_desc_sw128(abase + tap_off, p.desc_base_offset);
            const uint64_t bd0 = ptx::smem_desc_sw128(wbase, 0);
#pragma unroll
            for (int a = 0; a < C::APS; ++a) {
              if (a < a_lo || a >= a_hi) continue;
#pragma unroll
              for (int k = 0; k < KC / 16; ++k) {
                const uint64_t ad = ptx::desc_advance(ad0, a * (128 * 128) + k * 32);
                const uint64_t bd = ptx::desc_advance(bd0, k * 32);
                ptx::mma2_bf16_ss(tmem + (slot * C::APS + a) * BN, ad, bd, idesc, 1u);   // TMEM was pre-loaded with the bias
              }
            }
            ptx::tc_commit2_mc(w_empty + 8 * sw, (uint16_t)3);  // frees the stage in both CTAs
            if (++sw == C::SW) { sw = 0; pw ^= 1; }
          }
          ptx::tc_commit2_mc(a_empty + 8 * sa, (uint16_t)3);
          if (++sa == SA) { sa = 0; pa ^= 1; }
        }
        ptx::tc_commit2_mc(acc_full + 8 * slot, (uint16_t)3);
      }
      if (dbg && blockIdx.y == 0) {
        long long* d = p.dbg + (size_t)blockIdx.x * 8;
        d[0] = clock64() - t_begin; d[1] = t_acc; d[2] = t_a; d[3] = t_w;
      }
    }
  } else if (warp < ISS2_WARP) {
    // ============================== epilogue ==================================
    // 8 warps: TMEM lane quarter q = warp & 3 (hardware rule), column-group parity `par` = (warp-2) >> 2.
    // Units (accumulator a, 32-column group g) are software-pipelined: the tcgen05.ld (and the skip loads)
    // of unit u+1 are in flight while unit u is converted and stored.  The epilogue is instruction-issue
    // bound (measured), so: the folded bias lives IN the accumulator (each drained unit is immediately
    // re-initialised with tcgen05.st, MMAs always accumulate), ReLU is fused into cvt.rn.relu.bf16x2, and
    // the skip path is a template parameter.
    const int q = warp & 3, par = (warp - 2) >> 2;
    uint8_t* stage = sm + C::STAGE_OFF + (warp - 2) * 2048;
    constexpr int GPW = BN / 64;                        // column groups per warp per accumulator
    constexpr int UNITS = C::APS * GPW;                 // per slot
    const uint32_t tq = tmem + ((uint32_t)(q * 32) << 16);
    auto store_bias = [&](int a, int g) {
      uint32_t bv[32];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 t = *reinterpret_cast<const uint4*>(bias_s + g * 32 + j * 4);
        bv[4 * j] = t.x; bv[4 * j + 1] = t.y; bv[4 * j + 2] = t.z; bv[4 * j + 3] = t.w;
      }
      ptx::tmem_st_32x32(tq + a * BN + g * 32, bv);
    };
#pragma unroll
    for (int u = 0; u < 2 * UNITS; ++u) store_bias(u / GPW, (u % GPW) * 2 + par);
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncwarp();
    const uint32_t lead_empty = ptx::mapa(acc_empty, 0);   // the leader's barrier collects both CTAs' epilogue warps
    if (lane == 0) {                                    // completion 0 of both slots: accumulators hold the bias
      ptx::mbar_arrive_remote(lead_empty);
      ptx::mbar_arrive_remote(lead_empty + 8);
    }
    // Global accesses go through a 2 KB per-warp stage so that one instruction covers 8 pixels x 64 B
    // (lane l <-> pixel 8i + (l >> 2), 16-byte chunk l & 3) instead of 32 lines at a 256 B stride.
    // Stage layout: pixel px at px*64, chunk j at (j ^ ((px >> 1) & 3)) * 16: conflict-free both ways.
    const int cpx = lane >> 2, cch = lane & 3;
    int own_off[4], co_off[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      own_off[j] = lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4);
      const int px = 8 * j + cpx;
      co_off[j] = px * 64 + ((cch ^ ((px >> 1) & 3)) << 4);
    }
    for (int kk = 0; kk < iters; ++kk) {
      const uint32_t slot = kk & 1, pacc = (kk >> 1) & 1;
      const int it_raw = (int)blockIdx.x + kk * (int)gridDim.x;
      const bool live = it_raw < total;                 // a repeat item (pair padding) computes but does not store
      const int it = min(it_raw, total - 1);
      const int txi = it % p.tiles_x, r = it / p.tiles_x;
      const int tyi = r % p.tiles_y, n = r / p.tiles_y;
      const int x0 = txi * TILE_W;
      const int ybase = tyi * C::ROWS + q;              // accumulator row m = a*128 + q*32 + lane -> y = ybase + 4a
      // element offset of (pixel (ybase, x0 + cpx), channel nb*BN + cch*8); + 8*i pixels, + 4a rows, + g*32 channels
      const size_t e0 = (((size_t)n * p.H + ybase) * p.W + x0 + cpx) * p.Cout + nb * BN + cch * 8;
      const size_t acc_stride = (size_t)4 * p.W * p.Cout;
      const size_t px8 = (size_t)8 * p.Cout;
      bool pxok[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) pxok[i] = live && (8 * i + cpx < TILE_W) && (x0 + 8 * i + cpx < p.W);
      // DIRECT: this lane's own pixel (row ybase + 4a, column x0 + lane), channels nb*BN + g*32 ...
      const size_t eown = (((size_t)n * p.H + ybase) * p.W + x0 + lane) * p.Cout + nb * BN;
      const bool ownok = live && lane < TILE_W && x0 + lane < p.W;
      uint32_t v[2][32];
      uint4 sk[2][4];
      auto issue = [&](int u, int buf) {
        const int a = u / GPW, g = (u % GPW) * 2 + par;
        if (!(p.epi_variant & 2)) ptx::tmem_ld_32x32(tq + (slot * C::APS + a) * BN + g * 32, v[buf]);
        if (SKIP && !(p.epi_variant & 4)) {
          const bool yok = ybase + 4 * a < p.H;
          if (DIRECT) {
            const __nv_bfloat16* sp = p.skip + eown + a * acc_stride + g * 32;
#pragma unroll
            for (int i = 0; i < 4; ++i)
              sk[buf][i] = (yok && ownok) ? __ldg(reinterpret_cast<const uint4*>(sp + i * 8)) : make_uint4(0, 0, 0, 0);
          } else {
            const __nv_bfloat16* sp = p.skip + e0 + a * acc_stride + g * 32;
#pragma unroll
            for (int i = 0; i < 4; ++i)
              sk[buf][i] = (yok && pxok[i]) ? __ldg(reinterpret_cast<const uint4*>(sp + i * px8)) : make_uint4(0, 0, 0, 0);
          }
        }
      };
      const bool dbg = p.dbg != nullptr && warp == 2 && lane == 0 && blockIdx.y == 0;
      long long e0c = 0, e1c = 0;
      if (dbg) e0c = clock64();
      ptx::mbar_wait(acc_full + 8 * slot, pacc);
      if (dbg) { e1c = clock64(); p.dbg[(size_t)blockIdx.x * 8 + 4] += e1c - e0c; }
      ptx::tc_fence_after();
      issue(0, 0);
#pragma unroll
      for (int u = 0; u < UNITS; ++u) {
        const int buf = u & 1;
        const int a = u / GPW, g = (u % GPW) * 2 + par;
        ptx::tmem_ld_wait();
        store_bias(slot * C::APS + a, g);               // unit drained into registers: re-arm it for the next item
        if (u + 1 < UNITS) issue(u + 1, buf ^ 1);
        const bool yok = ybase + 4 * a < p.H;
        if (DIRECT) {
          uint4* op = reinterpret_cast<uint4*>(p.out + eown + a * acc_stride + g * 32);
          const bool ok = yok && ownok && !(p.epi_variant & 1);
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            uint32_t o[4];
            if (SKIP) {
              const uint32_t w[4] = {sk[buf][j4].x, sk[buf][j4].y, sk[buf][j4].z, sk[buf][j4].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float lo = fmaxf(__uint_as_float(v[buf][j4 * 8 + 2 * e]), 0.f) + ptx::bf16lo_to_f32(w[e]);
                const float hi = fmaxf(__uint_as_float(v[buf][j4 * 8 + 2 * e + 1]), 0.f) + ptx::bf16hi_to_f32(w[e]);
                o[e] = ptx::pack_bf16x2(lo, hi);
              }
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                o[e] = ptx::pack_relu_bf16x2(__uint_as_float(v[buf][j4 * 8 + 2 * e]), __uint_as_float(v[buf][j4 * 8 + 2 * e + 1]));
            }
            if (ok) op[j4] = make_uint4(o[0], o[1], o[2], o[3]);
          }
          continue;
        }
        if (SKIP) {
          uint4 sv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(stage + co_off[i]) = sk[buf][i];
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) sv[j] = *reinterpret_cast<const uint4*>(stage + own_off[j]);
          __syncwarp();
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {              // relu(acc) + x1, the skip AFTER the ReLU (app.py:292)
            const uint32_t w[4] = {sv[j4].x, sv[j4].y, sv[j4].z, sv[j4].w};
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float lo = fmaxf(__uint_as_float(v[buf][j4 * 8 + 2 * e]), 0.f) + ptx::bf16lo_to_f32(w[e]);
              const float hi = fmaxf(__uint_as_float(v[buf][j4 * 8 + 2 * e + 1]), 0.f) + ptx::bf16hi_to_f32(w[e]);
              o[e] = ptx::pack_bf16x2(lo, hi);
            }
            *reinterpret_cast<uint4*>(stage + own_off[j4]) = make_uint4(o[0], o[1], o[2], o[3]);
          }
        } else {
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e)
              o[e] = ptx::pack_relu_bf16x2(__uint_as_float(v[buf][j4 * 8 + 2 * e]), __uint_as_float(v[buf][j4 * 8 + 2 * e + 1]));
            *reinterpret_cast<uint4*>(stage + own_off[j4]) = make_uint4(o[0], o[1], o[2], o[3]);
          }
        }
        __syncwarp();
        {
          __nv_bfloat16* op = p.out + e0 + a * acc_stride + g * 32;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint4 o = *reinterpret_cast<const uint4*>(stage + co_off[i]);
            if (yok && pxok[i] && !(p.epi_variant & 1)) *reinterpret_cast<uint4*>(op + i * px8) = o;
          }
        }
        __syncwarp();
      }
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (dbg) p.dbg[(size_t)blockIdx.x * 8 + 5] += clock64() - e1c;
      if (lane == 0) ptx::mbar_arrive_remote(lead_empty + 8 * slot);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();                                  // no CTA leaves while its peer can still multicast to it
  if (warp == 1) {
    __syncwarp();
    ptx::tmem_dealloc2(tmem, 512);
  }
}

}  // namespace convtc3
}  // namespace vex
