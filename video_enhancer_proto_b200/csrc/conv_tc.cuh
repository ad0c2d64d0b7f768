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
// CTA = 6 warps: warp 0 TMA producer (1 lane), warp 1 MMA issuer (1 lane) + TMEM owner, warps 2-5 epilogue
// (TMEM lane quarter = warp & 3).  TMEM: all 512 columns = NACC accumulators of 128 pixels x BN channels.
// Persistent grid; work item = (image, 16- or 8-row x 30-column tile).
#pragma once

#include "ptx.cuh"

namespace vex {
namespace convtc {

constexpr int PITCH = 32;     // pixels per smem tile row (30 valid outputs + 2 halo)
constexpr int TILE_W = 30;
constexpr int KC = 64;        // channels per K chunk = one 128 B swizzle row
constexpr int SA = 2;         // activation stages
constexpr int NTHREADS = 192;

template <int BN>
struct Cfg {
  static_assert(BN == 128 || BN == 256, "BN");
  static constexpr int NACC = 512 / BN;                 // accumulators in TMEM
  static constexpr int ROWS = NACC * 4;                 // output rows per work item (128 px = 4 rows of 32)
  static constexpr int A_STAGE = (ROWS + 2) * PITCH * 128;
  static constexpr int W_STAGE = BN * 128;              // one (chunk, tap) weight slice: BN rows x 64 bf16
  static constexpr int SW = 4;
  static constexpr int BAR_OFF = SA * A_STAGE + SW * W_STAGE;
  static constexpr int BIAS_OFF = BAR_OFF + 256;
  static constexpr int SMEM = BIAS_OFF + BN * 4 + 1024; // + alignment slack
};

struct Params {
  const uint8_t* wpack;   // [cout_block][chunk][tap] slices, each BN x 128 B, pre-swizzled (SWIZZLE_128B image)
  const float* bias;      // [Cout] folded conv bias + BN shift
  __nv_bfloat16* out;     // NHWC [n][H][W][Cout]
  const __nv_bfloat16* skip;  // NHWC [n][H][W][Cout] added AFTER the ReLU (app.py:292), or nullptr
  int H, W, NC, Cout;     // NC = Cin / 64
  int tiles_x, tiles_y, n_img;
  int relu;
  int desc_base_offset;   // debug knob, see ptx::smem_desc_sw128
};

template <int BN>
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
  // barrier map (8 B each): a_full[SA] a_empty[SA] w_full[SW] w_empty[SW] acc_full acc_empty | tmem slot
  const uint32_t a_full = bar0, a_empty = bar0 + 8 * SA;
  const uint32_t w_full = a_empty + 8 * SA, w_empty = w_full + 8 * C::SW;
  const uint32_t acc_full = w_empty + 8 * C::SW, acc_empty = acc_full + 8;
  const uint32_t tmem_slot = acc_empty + 8;
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(sm + C::BAR_OFF + (tmem_slot - bar0));
  float* bias_s = reinterpret_cast<float*>(sm + C::BIAS_OFF);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nb = blockIdx.y;                            // Cout block
  const int total = p.tiles_x * p.tiles_y * p.n_img;

  if (threadIdx.x == 0) {
    for (int i = 0; i < SA; ++i) { ptx::mbar_init(a_full + 8 * i, 1); ptx::mbar_init(a_empty + 8 * i, 1); }
    for (int i = 0; i < C::SW; ++i) { ptx::mbar_init(w_full + 8 * i, 1); ptx::mbar_init(w_empty + 8 * i, 1); }
    ptx::mbar_init(acc_full, 1);
    ptx::mbar_init(acc_empty, 4);
    ptx::fence_mbar_init();
    ptx::prefetch_tensormap(&amap);
  }
  for (int i = threadIdx.x; i < BN; i += NTHREADS) bias_s[i] = p.bias[nb * BN + i];
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot_p;

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (lane == 0) {
      const uint8_t* wsrc = p.wpack + (size_t)nb * p.NC * 9 * C::W_STAGE;
      uint32_t sa = 0, pa = 0, sw = 0, pw = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x) {
        const int txi = it % p.tiles_x, r = it / p.tiles_x;
        const int tyi = r % p.tiles_y, n = r / p.tiles_y;
        const int x0 = txi * TILE_W, y0 = tyi * C::ROWS;
        for (int c = 0; c < p.NC; ++c) {
          ptx::mbar_wait(a_empty + 8 * sa, pa ^ 1);
          ptx::mbar_expect_tx(a_full + 8 * sa, C::A_STAGE);
          ptx::tma_load_4d(a_stage0 + sa * C::A_STAGE, &amap, a_full + 8 * sa, c * KC, x0 - 1, y0 - 1, n);
          if (++sa == SA) { sa = 0; pa ^= 1; }
          for (int t = 0; t < 9; ++t) {
            ptx::mbar_wait(w_empty + 8 * sw, pw ^ 1);
            ptx::mbar_expect_tx(w_full + 8 * sw, C::W_STAGE);
            ptx::bulk_load(w_stage0 + sw * C::W_STAGE, wsrc + (size_t)(c * 9 + t) * C::W_STAGE, C::W_STAGE,
                           w_full + 8 * sw);
            if (++sw == C::SW) { sw = 0; pw ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer ================================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::idesc_bf16_f32(128, BN);
      uint32_t sa = 0, pa = 0, sw = 0, pw = 0, pacc = 0;
      for (int it = blockIdx.x; it < total; it += gridDim.x) {
        ptx::mbar_wait(acc_empty, pacc ^ 1);            // epilogue has drained the previous item
        ptx::tc_fence_after();
        for (int c = 0; c < p.NC; ++c) {
          ptx::mbar_wait(a_full + 8 * sa, pa);
          const uint32_t abase = a_stage0 + sa * C::A_STAGE;
          for (int t = 0; t < 9; ++t) {
            ptx::mbar_wait(w_full + 8 * sw, pw);
            ptx::tc_fence_after();
            const uint32_t wbase = w_stage0 + sw * C::W_STAGE;
            const uint32_t tap_off = ((t / 3) * PITCH + (t % 3)) * 128;
#pragma unroll
            for (int a = 0; a < C::NACC; ++a) {
#pragma unroll
              for (int k = 0; k < KC / 16; ++k) {
                const uint64_t ad = ptx::smem_desc_sw128(abase + tap_off + a * (128 * 128) + k * 32,
                                                         p.desc_base_offset);
                const uint64_t bd = ptx::smem_desc_sw128(wbase + k * 32, 0);
                ptx::mma_bf16_ss(tmem + a * BN, ad, bd, idesc, (c | t | k) != 0);
              }
            }
            ptx::tc_commit(w_empty + 8 * sw);           // slice consumed once these MMAs retire
            if (++sw == C::SW) { sw = 0; pw ^= 1; }
          }
          ptx::tc_commit(a_empty + 8 * sa);
          if (++sa == SA) { sa = 0; pa ^= 1; }
        }
        ptx::tc_commit(acc_full);
        pacc ^= 1;
      }
    }
  } else {
    // ============================== epilogue ==================================
    const int q = warp & 3;                             // TMEM lane quarter this warp may read
    uint32_t pacc = 0;
    for (int it = blockIdx.x; it < total; it += gridDim.x) {
      const int txi = it % p.tiles_x, r = it / p.tiles_x;
      const int tyi = r % p.tiles_y, n = r / p.tiles_y;
      const int x = txi * TILE_W + lane;
      ptx::mbar_wait(acc_full, pacc);
      pacc ^= 1;
      ptx::tc_fence_after();
#pragma unroll 1
      for (int a = 0; a < C::NACC; ++a) {
        const int y = tyi * C::ROWS + a * 4 + q;        // accumulator row m = a*128 + q*32 + lane
        const bool valid = (lane < TILE_W) && (x < p.W) && (y < p.H);
        const size_t pix = ((size_t)n * p.H + y) * p.W + x;
        __nv_bfloat16* optr = p.out + pix * p.Cout + nb * BN;
        const __nv_bfloat16* sptr = p.skip ? p.skip + pix * p.Cout + nb * BN : nullptr;
#pragma unroll 1
        for (int g = 0; g < BN / 32; ++g) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + a * BN + g * 32, v);
          ptx::tmem_ld_wait();
          if (valid) {
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              f[j] = __uint_as_float(v[j]) + bias_s[g * 32 + j];
              if (p.relu) f[j] = fmaxf(f[j], 0.0f);
            }
            if (sptr) {
              const uint4* s4 = reinterpret_cast<const uint4*>(sptr + g * 32);
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4) {
                const uint4 s = __ldg(s4 + j4);
                const uint32_t w[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  f[j4 * 8 + 2 * e] += ptx::bf16lo_to_f32(w[e]);
                  f[j4 * 8 + 2 * e + 1] += ptx::bf16hi_to_f32(w[e]);
                }
              }
            }
            uint4* o4 = reinterpret_cast<uint4*>(optr + g * 32);
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              uint4 o;
              o.x = ptx::pack_bf16x2(f[j4 * 8 + 0], f[j4 * 8 + 1]);
              o.y = ptx::pack_bf16x2(f[j4 * 8 + 2], f[j4 * 8 + 3]);
              o.z = ptx::pack_bf16x2(f[j4 * 8 + 4], f[j4 * 8 + 5]);
              o.w = ptx::pack_bf16x2(f[j4 * 8 + 6], f[j4 * 8 + 7]);
              o4[j4] = o;
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc_empty);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    ptx::tmem_dealloc(tmem, 512);
  }
}

}  // namespace convtc
}  // namespace vex
