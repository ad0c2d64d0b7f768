// AdvancedUpscaler (reference app.py:269-301) on sm_100a: weight folding/packing, workspace, layer launches.
//
//   u8 HWC --conv1_u8--> x1 --conv_tc(+x1)--> x2 --conv_tc--> x3 --conv_tc--> x4 --conv3x3_to3--> x5
//          --upconv1--> up (2H x 2W x up_ch) --conv3x3_to3<clamp,u8>--> u8 HWC (2H, 2W)
//
// Activations x1..x4 and `up` are NHWC bf16 in the ctx workspace; x5 is fp32.  conv2..conv4 (97.8 % of the
// FLOPs) run on tcgen05 (conv_tc.cuh); the thin layers are fp32 CUDA-core kernels.
#include <algorithm>
#include <math.h>
#include <string.h>

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <nvtx3/nvToolsExt.h>

#include "cnn_small.cuh"
#include "conv_tc.cuh"
#include "conv_tc2.cuh"
#include "conv_tc3.cuh"
#include "thin_tc.cuh"
#include "tail_tc.cuh"
#include "tail_tc2.cuh"
#include "tail_tc3.cuh"
#include "vex_internal.cuh"

namespace vex {

struct TcLayer {
  int cin = 0, cout = 0, bn = 128;       // bn = Cout block per CTA (TMEM accumulator width)
  uint8_t* wpack = nullptr;              // device: [cout/bn][cin/64][9] swizzled slices
  float* bias = nullptr;                 // device: [cout]
  std::vector<float> w_folded;           // host fp32 [cout][cin][3][3] * bn scale (kept for re-packing)
};

struct CnnPlan {
  vex_cnn_desc arch{};
  // thin layers (device, fp32)
  float *w1g = nullptr, *b1 = nullptr;       // conv1: [27][c1], [c1]
  float *w5g = nullptr, *b5 = nullptr;       // conv5: [c4/64][9][64][4], [3]
  float *wu1 = nullptr, *bu1 = nullptr;      // upconv1: [4][12][up], [up]
  float *wu2 = nullptr, *bu2 = nullptr;      // upconv2: [up/64][9][64][4], [3]
  TcLayer tc[3];                             // conv2, conv3, conv4
  TcLayer extra[VEX_CNN_MAX_EXTRA];          // deepened plans: c3 -> c3 blocks after conv3
  // thin layers on the tensor cores (thin_tc.cuh): SWIZZLE_128B operand images, bf16
  uint8_t *t1w = nullptr, *tu1wb = nullptr, *t5w = nullptr, *tu2w = nullptr;   // t1w / tu1wb carry their bias in k = 27, 28
  uint8_t* tu2w3 = nullptr;                  // upconv2 for tail_tc3.cuh: [dy] slices of 16 rows (dx*3 + co) x 64 k
  // workspace
  uint8_t* ws = nullptr;
  size_t ws_bytes = 0;
  void* encode = nullptr;                    // cuTensorMapEncodeTiled, resolved through the runtime
  uint8_t* t1wh[2] = {nullptr, nullptr};     // conv1, fp16 offset-1024 form (MODE_CONV1_H) for RGB / BGR input buffers
  uint8_t* t1wf = nullptr;                   // conv1 for fp32 input (vex_cnn2x_f32in): no 1/255 fold, weights repeated at k = 32..58 for the lo halves
  std::vector<cudaEvent_t> prof;             // option "profile": 8 events per sub-batch of the last forward
  bool prof_folded = false;                  // the last forward is already inside the statistics
  std::vector<std::vector<cudaEvent_t>> prof_done;   // earlier forwards not yet folded into the statistics
  std::vector<cudaEvent_t> prof_pool;
  double prof_sum[7] = {0, 0, 0, 0, 0, 0, 0};
  float prof_min[7] = {0, 0, 0, 0, 0, 0, 0}, prof_max[7] = {0, 0, 0, 0, 0, 0, 0};
  int prof_n = 0;
  // vex_cnn2x_u8_noalloc: this device's address of g_noalloc_ws (resolved once, at weight load), or null + the reason
  uint8_t* noalloc_ws = nullptr;
  cudaError_t noalloc_err = cudaSuccess;
};

// Scratch of the no-allocation entry point (vex_cnn2x_u8_noalloc, include/vex.h): the activation workspace of the CNN for
// callers that may not allocate (CUDA graph capture, serving loops).  File-scope device storage: every GPU whose context
// loads this library holds VEX_NOALLOC_SCRATCH_BYTES of it until the process exits.  In HOST code this name is only a
// stand-in for the symbol; the device address comes from cudaGetSymbolAddress (cnn_noalloc_setup), once per device.
static __device__ __align__(256) unsigned char g_noalloc_ws[VEX_NOALLOC_SCRATCH_BYTES];

namespace {

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;      // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

template <typename T>
int upload(vex_ctx* ctx, T** dst, const void* src, size_t bytes) {
  if (*dst) cudaFree(*dst);
  *dst = nullptr;
  VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(dst), bytes));
  VEX_CUDA(ctx, cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice));
  return VEX_OK;
}

// [cout/bn][cin/64][9] slices; slice = bn rows x 64 bf16 (128 B), stored as the SWIZZLE_128B image of a
// K-major operand: 16-byte chunk j of row r lives at chunk position j ^ (r & 7).
int pack_tc(vex_ctx* ctx, TcLayer& L) {
  const int nb = L.cout / L.bn, nc = L.cin / 64;
  std::vector<uint16_t> img((size_t)L.cout * L.cin * 9);
  for (int b = 0; b < nb; ++b)
    for (int c = 0; c < nc; ++c)
      for (int t = 0; t < 9; ++t) {
        uint16_t* slice = img.data() + ((size_t)(b * nc + c) * 9 + t) * L.bn * 64;
        for (int r = 0; r < L.bn; ++r)
          for (int k = 0; k < 64; ++k) {
            const int co = b * L.bn + r, ci = c * 64 + k;
            const float v = L.w_folded[((size_t)co * L.cin + ci) * 9 + t];
            const int chunk = (k >> 3) ^ (r & 7);
            slice[(size_t)r * 64 + chunk * 8 + (k & 7)] = f32_to_bf16_rne(v);
          }
      }
  return upload(ctx, &L.wpack, img.data(), img.size() * 2);
}

int make_amap(vex_ctx* ctx, CUtensorMap* map, const void* base, int n, int H, int W, int C, int box_rows) {
  CnnPlan* P = ctx->cnn;
  const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
  const cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  const cuuint32_t box[4] = {64, (cuuint32_t)convtc::PITCH, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = reinterpret_cast<EncodeFn>(P->encode)(
      map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(ctx, VEX_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  return VEX_OK;
}

template <int BN, bool SKIP>
int launch_tc(vex_ctx* ctx, const TcLayer& L, const __nv_bfloat16* in, __nv_bfloat16* out,
              const __nv_bfloat16* skip, int n, int H, int W, cudaStream_t s) {
  using C = convtc::Cfg<BN>;
  CUtensorMap amap;
  int rc = make_amap(ctx, &amap, in, n, H, W, L.cin, C::ROWS + 2);
  if (rc) return rc;
  convtc::Params p;
  p.wpack = L.wpack; p.bias = L.bias; p.out = out; p.skip = skip;
  p.H = H; p.W = W; p.NC = L.cin / 64; p.Cout = L.cout;
  p.tiles_x = (W + convtc::TILE_W - 1) / convtc::TILE_W;
  p.tiles_y = (H + C::ROWS - 1) / C::ROWS;
  p.n_img = n; p.desc_base_offset = ctx->opt_desc_base_offset;
  p.dbg = nullptr;
  p.epi_variant = ctx->opt_epi_variant;
  if (ctx->opt_conv_dbg) {                   // bring-up: per-CTA cycle counters, printed after the launch
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&p.dbg), (size_t)ctx->sm_count * 64));
    VEX_CUDA(ctx, cudaMemsetAsync(p.dbg, 0, (size_t)ctx->sm_count * 64, s));
  }
  const long long total = (long long)p.tiles_x * p.tiles_y * n;
  if (total > 0x7fffffffLL) return fail(ctx, VEX_ERR_INVALID, "conv_tc: too many tiles");
  const unsigned abit = (BN == 256 ? 2u : 1u) << (SKIP ? 8 : 0);
  if (!(ctx->attr_mask & abit)) {            // per device, hence per ctx
    VEX_CUDA(ctx, cudaFuncSetAttribute(convtc::conv3x3_tc_kernel<BN, SKIP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       C::SMEM));
    ctx->attr_mask |= abit;
  }
  dim3 grid((unsigned)std::min<long long>(total, ctx->sm_count), L.cout / BN);
  convtc::conv3x3_tc_kernel<BN, SKIP><<<grid, convtc::NTHREADS, C::SMEM, s>>>(amap, p);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  if (p.dbg) {
    std::vector<long long> h((size_t)ctx->sm_count * 8);
    VEX_CUDA(ctx, cudaStreamSynchronize(s));
    VEX_CUDA(ctx, cudaMemcpy(h.data(), p.dbg, h.size() * 8, cudaMemcpyDeviceToHost));
    cudaFree(p.dbg);
    double a[6] = {0, 0, 0, 0, 0, 0};
    for (unsigned b = 0; b < grid.x; ++b)
      for (int i = 0; i < 6; ++i) a[i] += (double)h[(size_t)b * 8 + i] / grid.x;
    const double items = (double)total / grid.x;
    fprintf(stderr, "[conv_dbg] cin=%d cout=%d bn=%d items/cta=%.1f cycles/item: total=%.0f mma_wait_acc=%.0f "
                    "mma_wait_a=%.0f mma_wait_w=%.0f | epi_wait_full=%.0f epi_work=%.0f\n",
            L.cin, L.cout, BN, items, a[0] / items, a[1] / items, a[2] / items, a[3] / items, a[4] / items,
            a[5] / items);
  }
  return VEX_OK;
}

// v2: cluster-of-two kernel with double-buffered TMEM slots and multicast weights (conv_tc2.cuh)
template <int BN, bool SKIP, bool DIRECT>
int launch_tc2(vex_ctx* ctx, const TcLayer& L, const __nv_bfloat16* in, __nv_bfloat16* out,
               const __nv_bfloat16* skip, int n, int H, int W, cudaStream_t s) {
  using C = convtc2::Cfg<BN>;
  CUtensorMap amap;
  int rc = make_amap(ctx, &amap, in, n, H, W, L.cin, C::ROWS + 2);
  if (rc) return rc;
  convtc2::Params p;
  p.wpack = L.wpack; p.bias = L.bias; p.out = out; p.skip = skip;
  p.H = H; p.W = W; p.NC = L.cin / 64; p.Cout = L.cout;
  p.tiles_x = (W + convtc2::TILE_W - 1) / convtc2::TILE_W;
  p.tiles_y = (H + C::ROWS - 1) / C::ROWS;
  p.n_img = n; p.desc_base_offset = ctx->opt_desc_base_offset;
  p.dbg = nullptr;
  p.epi_variant = ctx->opt_epi_variant;
  if (ctx->opt_conv_dbg) {
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&p.dbg), (size_t)ctx->sm_count * 64));
    VEX_CUDA(ctx, cudaMemsetAsync(p.dbg, 0, (size_t)ctx->sm_count * 64, s));
  }
  const long long total = (long long)p.tiles_x * p.tiles_y * n;
  if (total > 0x3fffffffLL) return fail(ctx, VEX_ERR_INVALID, "conv_tc2: too many tiles");
  auto kern = convtc2::conv3x3_tc_kernel<BN, SKIP, DIRECT>;
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(convtc2::NTHREADS);
  cfg.dynamicSmemBytes = C::SMEM;
  cfg.stream = s;
  cfg.attrs = attr; cfg.numAttrs = 1;
  const unsigned abit = ((BN == 256 ? 2u : 1u) << (SKIP ? 8 : 0)) << (DIRECT ? 20 : 16);
  if (!(ctx->attr_mask & abit)) {            // per device, hence per ctx
    VEX_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    ctx->attr_mask |= abit;
  }
  if (ctx->max_pairs <= 0) {                 // how many 2-CTA clusters of this footprint are co-resident
    cfg.gridDim = dim3(2 * (unsigned)(ctx->sm_count / 2), 1);
    int nc = 0;
    VEX_CUDA(ctx, cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
    ctx->max_pairs = std::max(1, std::min(nc, ctx->sm_count / 2));
  }
  const int nby = L.cout / BN;
  const long long pairs = std::min<long long>((total + 1) / 2, std::max(1, ctx->max_pairs / nby));
  cfg.gridDim = dim3((unsigned)(2 * pairs), nby);
  VEX_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern, amap, p));
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  if (p.dbg) {
    std::vector<long long> h((size_t)ctx->sm_count * 8);
    VEX_CUDA(ctx, cudaStreamSynchronize(s));
    VEX_CUDA(ctx, cudaMemcpy(h.data(), p.dbg, h.size() * 8, cudaMemcpyDeviceToHost));
    cudaFree(p.dbg);
    double a[6] = {0, 0, 0, 0, 0, 0};
    for (unsigned b = 0; b < cfg.gridDim.x; ++b)
      for (int i = 0; i < 6; ++i) a[i] += (double)h[(size_t)b * 8 + i] / cfg.gridDim.x;
    const double items = (double)total / cfg.gridDim.x;
    fprintf(stderr, "[conv_dbg v2] cin=%d cout=%d bn=%d pairs=%lld items/cta=%.1f cycles/item: total=%.0f mma_wait_acc=%.0f "
                    "mma_wait_a=%.0f mma_wait_w=%.0f | epi_wait_full=%.0f epi_work=%.0f\n",
            L.cin, L.cout, BN, pairs, items, a[0] / items, a[1] / items, a[2] / items, a[3] / items, a[4] / items,
            a[5] / items);
  }
  return VEX_OK;
}

// v3: cluster-of-two kernel issuing cta_group::2 MMAs (M = 256 over the pair, B split across the pair; conv_tc3.cuh)
template <int BN, bool SKIP, bool DIRECT>
int launch_tc3(vex_ctx* ctx, const TcLayer& L, const __nv_bfloat16* in, __nv_bfloat16* out,
               const __nv_bfloat16* skip, int n, int H, int W, cudaStream_t s) {
  using C = convtc3::Cfg<BN>;
  CUtensorMap amap;
  int rc = make_amap(ctx, &amap, in, n, H, W, L.cin, C::ROWS + 2);
  if (rc) return rc;
  convtc3::Params p;
  p.wpack = L.wpack; p.bias = L.bias; p.out = out; p.skip = skip;
  p.H = H; p.W = W; p.NC = L.cin / 64; p.Cout = L.cout;
  p.tiles_x = (W + convtc3::TILE_W - 1) / convtc3::TILE_W;
  p.tiles_y = (H + C::ROWS - 1) / C::ROWS;
  p.n_img = n; p.desc_base_offset = ctx->opt_desc_base_offset;
  p.dbg = nullptr;
  p.epi_variant = ctx->opt_epi_variant;
  if (ctx->opt_conv_dbg) {
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&p.dbg), (size_t)ctx->sm_count * 64));
    VEX_CUDA(ctx, cudaMemsetAsync(p.dbg, 0, (size_t)ctx->sm_count * 64, s));
  }
  const long long total = (long long)p.tiles_x * p.tiles_y * n;
  if (total > 0x3fffffffLL) return fail(ctx, VEX_ERR_INVALID, "conv_tc3: too many tiles");
  auto kern = convtc3::conv3x3_tc_kernel<BN, SKIP, DIRECT>;
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(convtc3::NTHREADS);
  cfg.dynamicSmemBytes = C::SMEM;
  cfg.stream = s;
  cfg.attrs = attr; cfg.numAttrs = 1;
  const unsigned abit = ((BN == 256 ? 2u : 1u) << (SKIP ? 8 : 0)) << (DIRECT ? 20 : 16);
  if (!(ctx->attr_mask3 & abit)) {            // per device, hence per ctx
    VEX_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    ctx->attr_mask3 |= abit;
  }
  if (ctx->max_pairs3 <= 0) {                 // how many 2-CTA clusters of this footprint are co-resident
    cfg.gridDim = dim3(2 * (unsigned)(ctx->sm_count / 2), 1);
    int nc = 0;
    VEX_CUDA(ctx, cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
    ctx->max_pairs3 = std::max(1, std::min(nc, ctx->sm_count / 2));
  }
  const int nby = L.cout / BN;
  const long long pairs = std::min<long long>((total + 1) / 2, std::max(1, ctx->max_pairs3 / nby));
  cfg.gridDim = dim3((unsigned)(2 * pairs), nby);
  VEX_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern, amap, p));
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  if (p.dbg) {
    std::vector<long long> h((size_t)ctx->sm_count * 8);
    VEX_CUDA(ctx, cudaStreamSynchronize(s));
    VEX_CUDA(ctx, cudaMemcpy(h.data(), p.dbg, h.size() * 8, cudaMemcpyDeviceToHost));
    cudaFree(p.dbg);
    double a[6] = {0, 0, 0, 0, 0, 0};
    for (unsigned b = 0; b < cfg.gridDim.x; ++b)
      for (int i = 0; i < 6; ++i) a[i] += (double)h[(size_t)b * 8 + i] / cfg.gridDim.x;
    const double items = (double)total / cfg.gridDim.x;
    fprintf(stderr, "[conv_dbg v3] cin=%d cout=%d bn=%d pairs=%lld items/cta=%.1f cycles/item: total=%.0f mma_wait_acc=%.0f "
                    "mma_wait_a=%.0f mma_wait_w=%.0f | epi_wait_full=%.0f epi_work=%.0f\n",
            L.cin, L.cout, BN, pairs, items, a[0] / items, a[1] / items, a[2] / items, a[3] / items, a[4] / items,
            a[5] / items);
  }
  return VEX_OK;
}

#ifndef VEX_CONV_DEFAULT_VARIANT
#define VEX_CONV_DEFAULT_VARIANT 3
#endif

int run_tc(vex_ctx* ctx, const TcLayer& L, const __nv_bfloat16* in, __nv_bfloat16* out,
           const __nv_bfloat16* skip, int n, int H, int W, cudaStream_t s) {
  // conv_variant: 1 = v1 single-CTA kernel; 2 = v2 cluster pairs with multicast weights (cta_group::1);
  // 3 = v3 cta_group::2 pairs; 0 = auto (see below)
  const int variant = ctx->opt_conv_variant == 0 ? VEX_CONV_DEFAULT_VARIANT : ctx->opt_conv_variant;
  if (variant == 3) {
    const bool direct = (ctx->opt_epi_variant & 8) != 0;
    if (direct) {
      if (skip)
        return L.bn == 256 ? launch_tc3<256, true, true>(ctx, L, in, out, skip, n, H, W, s)
                           : launch_tc3<128, true, true>(ctx, L, in, out, skip, n, H, W, s);
      return L.bn == 256 ? launch_tc3<256, false, true>(ctx, L, in, out, skip, n, H, W, s)
                         : launch_tc3<128, false, true>(ctx, L, in, out, skip, n, H, W, s);
    }
    if (skip)
      return L.bn == 256 ? launch_tc3<256, true, false>(ctx, L, in, out, skip, n, H, W, s)
                         : launch_tc3<128, true, false>(ctx, L, in, out, skip, n, H, W, s);
    return L.bn == 256 ? launch_tc3<256, false, false>(ctx, L, in, out, skip, n, H, W, s)
                       : launch_tc3<128, false, false>(ctx, L, in, out, skip, n, H, W, s);
  }
  const bool v2 = variant == 2;
  if (v2) {
    const bool direct = (ctx->opt_epi_variant & 8) != 0;
    if (direct) {
      if (skip)
        return L.bn == 256 ? launch_tc2<256, true, true>(ctx, L, in, out, skip, n, H, W, s)
                           : launch_tc2<128, true, true>(ctx, L, in, out, skip, n, H, W, s);
      return L.bn == 256 ? launch_tc2<256, false, true>(ctx, L, in, out, skip, n, H, W, s)
                         : launch_tc2<128, false, true>(ctx, L, in, out, skip, n, H, W, s);
    }
    if (skip)
      return L.bn == 256 ? launch_tc2<256, true, false>(ctx, L, in, out, skip, n, H, W, s)
                         : launch_tc2<128, true, false>(ctx, L, in, out, skip, n, H, W, s);
    return L.bn == 256 ? launch_tc2<256, false, false>(ctx, L, in, out, skip, n, H, W, s)
                       : launch_tc2<128, false, false>(ctx, L, in, out, skip, n, H, W, s);
  }
  if (skip)
    return L.bn == 256 ? launch_tc<256, true>(ctx, L, in, out, skip, n, H, W, s)
                       : launch_tc<128, true>(ctx, L, in, out, skip, n, H, W, s);
  return L.bn == 256 ? launch_tc<256, false>(ctx, L, in, out, skip, n, H, W, s)
                     : launch_tc<128, false>(ctx, L, in, out, skip, n, H, W, s);
}

// rows x (nchunk x 64) bf16 matrix -> [nchunk] slices of rows x 128 B in the SWIZZLE_128B K-major image
template <typename F>
int pack_rows_sw128(vex_ctx* ctx, uint8_t** dst, int rows, int nchunk, F value) {
  std::vector<uint16_t> img((size_t)nchunk * rows * 64, 0);
  for (int c = 0; c < nchunk; ++c)
    for (int r = 0; r < rows; ++r)
      for (int k = 0; k < 64; ++k)
        img[((size_t)c * rows + r) * 64 + (((k >> 3) ^ (r & 7)) << 3) + (k & 7)] = f32_to_bf16_rne(value(r, c * 64 + k));
  return upload(ctx, dst, img.data(), img.size() * 2);
}

template <int MODE, int BN>
int launch_i27(vex_ctx* ctx, unsigned abit, const void* in, const uint8_t* w, __nv_bfloat16* out,
               int n, int H, int W, int in_bgr, cudaStream_t s) {
  using C = thin::I27Cfg<BN>;
  if (!(ctx->attr_mask & abit)) {
    VEX_CUDA(ctx, cudaFuncSetAttribute(thin::im2col27_kernel<MODE, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       C::SMEM));
    ctx->attr_mask |= abit;
  }
  thin::I27Params p;
  p.in = in; p.wpack = w; p.out = out; p.H = H; p.W = W;   // the bias travels inside w (k = 27, 28)
  p.npix = (long long)n * H * W; p.in_bgr = in_bgr; p.probe = ctx->opt_thin_probe;
  p.out_cs = BN; p.dbg = nullptr;
  const long long items = (p.npix + 127) / 128;
  const unsigned grid = (unsigned)std::min<long long>(items, ctx->sm_count);
  if (ctx->opt_conv_dbg) {
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&p.dbg), (size_t)grid * 64));
    VEX_CUDA(ctx, cudaMemsetAsync(p.dbg, 0, (size_t)grid * 64, s));
  }
  thin::im2col27_kernel<MODE, BN><<<grid, thin::I27_THREADS, C::SMEM, s>>>(p);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  if (p.dbg) {
    std::vector<long long> h((size_t)grid * 8);
    VEX_CUDA(ctx, cudaStreamSynchronize(s));
    VEX_CUDA(ctx, cudaMemcpy(h.data(), p.dbg, h.size() * 8, cudaMemcpyDeviceToHost));
    cudaFree(p.dbg);
    double a[8] = {0};
    for (unsigned b = 0; b < grid; ++b)
      for (int i = 0; i < 8; ++i) a[i] += (double)h[(size_t)b * 8 + i] / grid;
    const double it = (double)items / grid;
    fprintf(stderr, "[im2col27_dbg mode=%d bn=%d] items/cta=%.1f cycles/item: MMA total=%.0f wait_acc_empty=%.0f wait_a_full=%.0f | "
                    "builder(group 0, per own item) total=%.0f wait_a_empty=%.0f | drain(group 0, per own item) total=%.0f "
                    "wait_acc_full=%.0f wait_bulk_read=%.0f\n",
            MODE, BN, it, a[0] / it, a[1] / it, a[2] / it, 2 * a[4] / it, 2 * a[3] / it, 2 * a[7] / it, 2 * a[5] / it, 2 * a[6] / it);
  }
  return VEX_OK;
}

template <int MODE>
int launch_g27(vex_ctx* ctx, unsigned abit, const __nv_bfloat16* in, int C, const uint8_t* w, const float* bias,
               float* out_f4, uint8_t* out_u8, float* out_nchw, int n, int H, int W, int out_bgr, cudaStream_t s) {
  if (!(ctx->attr_mask & abit)) {
    VEX_CUDA(ctx, cudaFuncSetAttribute(thin::gemm27_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       thin::G27Cfg::SMEM));
    ctx->attr_mask |= abit;
  }
  CUtensorMap amap;
  int rc = make_amap(ctx, &amap, in, n, H, W, C, 8);
  if (rc) return rc;
  thin::G27Params p;
  p.wpack = w; p.bias = bias; p.out_f4 = reinterpret_cast<float4*>(out_f4); p.out_u8 = out_u8; p.out_nchw = out_nchw;
  p.H = H; p.W = W; p.NC = C / 64;
  p.tiles_x = (W + thin::G27_TW - 1) / thin::G27_TW;
  p.tiles_y = (H + thin::G27_ROWS - 1) / thin::G27_ROWS;
  p.n_img = n; p.out_bgr = out_bgr;
  const long long total = (long long)p.tiles_x * p.tiles_y * n;
  if (total > 0x7fffffffLL) return fail(ctx, VEX_ERR_INVALID, "gemm27: too many tiles");
  thin::gemm27_kernel<MODE><<<(unsigned)std::min<long long>(total, ctx->sm_count), thin::G27_THREADS,
                              thin::G27Cfg::SMEM, s>>>(amap, p);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  return VEX_OK;
}

// x5 [n][H][W] -> uint8 [n][2H][2W][3]: upconv1 + relu + upconv2 + clamp + quantise in one kernel (tail_tc.cuh)
int launch_uptail(vex_ctx* ctx, const float* x5, const uint8_t* w1, const uint8_t* w2, const float* b2,
                  uint8_t* out_u8, float* out_nchw, int n, int H, int W, int out_bgr, cudaStream_t s) {
  // fuse_tail: 4 = default (tail_tc3.cuh, below); 1 = the first fused tail (relu(upconv1) handed to the second GEMM through shared memory, tail_tc.cuh); 3 = the variant
  // that keeps it in tensor memory (tail_tc2.cuh: bit-identical, 40 % less shared-memory traffic, two issuing threads; it
  // measures 2007-2022 against 1960 cycles per item - a tie); 4 = tail_tc3.cuh (vertical tap of upconv2 in the tensor core,
  // horizontal tap by lane shuffles, 10 x 30 outputs per item from a 7 x 17 low-res patch: 2.7-2.9 ms against 3.4-3.9)
  const bool v2 = ctx->opt_fuse_tail == 3, v3 = ctx->opt_fuse_tail == 4;   // 4: tail_tc3.cuh (vertical tap in the tensor core)
  if (!(ctx->attr_mask & 0x400u)) {
    VEX_CUDA(ctx, cudaFuncSetAttribute(thin::uptail3_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       thin::UT3Cfg::SMEM));
    VEX_CUDA(ctx, cudaFuncSetAttribute(thin::uptail3_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       thin::UT3Cfg::SMEM));
    VEX_CUDA(ctx, cudaFuncSetAttribute(thin::uptail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       thin::UTCfg::SMEM));
    VEX_CUDA(ctx, cudaFuncSetAttribute(thin::uptail2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       thin::UT2Cfg::SMEM));
    ctx->attr_mask |= 0x400u;
  }
  thin::UTParams p;
  p.x5 = reinterpret_cast<const float4*>(x5); p.w1pack = w1; p.w2pack = w2; p.bias2 = b2;
  p.out_u8 = out_u8; p.out_nchw = out_nchw; p.H = H; p.W = W;
  p.tiles_x = (2 * W + thin::G27_TW - 1) / thin::G27_TW;
  p.tiles_y = v3 ? (2 * H + thin::UT3_OROWS - 1) / thin::UT3_OROWS : (2 * H + thin::G27_ROWS - 1) / thin::G27_ROWS;
  p.n_img = n; p.out_bgr = out_bgr; p.probe = ctx->opt_thin_probe;
  p.dbg = nullptr;
  const long long total = (long long)p.tiles_x * p.tiles_y * n;
  if (total > 0x7fffffffLL) return fail(ctx, VEX_ERR_INVALID, "uptail: too many tiles");
  p.rcp_tx = p.rcp_ty = 0.f;
  if (total < (1LL << 22)) { p.rcp_tx = 1.0f / (float)p.tiles_x; p.rcp_ty = 1.0f / (float)p.tiles_y; }
  const unsigned grid = (unsigned)std::min<long long>(total, ctx->sm_count);
  if (ctx->opt_conv_dbg) {                   // bring-up: per-role cycle counters, printed after the launch
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&p.dbg), (size_t)grid * 128));
    VEX_CUDA(ctx, cudaMemsetAsync(p.dbg, 0, (size_t)grid * 128, s));
  }
  CUtensorMap xmap;                          // x5 as [n][H][W][4] fp32: box = the 19 x 7 pixel window of one item, OOB = 0
  {
    CnnPlan* P = ctx->cnn;
    const cuuint64_t dims[4] = {4, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
    const cuuint64_t strides[3] = {16, (cuuint64_t)W * 16, (cuuint64_t)H * W * 16};
    const cuuint32_t box[4] = {4, (cuuint32_t)(v3 ? thin::UT3_XW : thin::UT_XW), (cuuint32_t)(v3 ? thin::UT3_XH : thin::UT_XH), 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = reinterpret_cast<EncodeFn>(P->encode)(
        &xmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x5), dims, strides, box, estr,
        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ctx, VEX_ERR_CUDA, "cuTensorMapEncodeTiled (x5) failed (" + std::to_string((int)r) + ")");
  }
  // the production instantiation carries neither the cycle counters nor the fp32 output of the parity hook
  if (v3 && (p.dbg || out_nchw)) thin::uptail3_kernel<true, true><<<grid, thin::UT3_THREADS, thin::UT3Cfg::SMEM, s>>>(xmap, p);
  else if (v3) thin::uptail3_kernel<false, false><<<grid, thin::UT3_THREADS, thin::UT3Cfg::SMEM, s>>>(xmap, p);
  else if (v2) thin::uptail2_kernel<<<grid, thin::UT_THREADS, thin::UT2Cfg::SMEM, s>>>(xmap, p);
  else thin::uptail_kernel<<<grid, thin::UT_THREADS, thin::UTCfg::SMEM, s>>>(xmap, p);
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  if (p.dbg) {
    std::vector<long long> h((size_t)grid * 16);
    VEX_CUDA(ctx, cudaStreamSynchronize(s));
    VEX_CUDA(ctx, cudaMemcpy(h.data(), p.dbg, h.size() * 8, cudaMemcpyDeviceToHost));
    cudaFree(p.dbg);
    double a[16] = {0};
    for (unsigned b = 0; b < grid; ++b)
      for (int i = 0; i < 16; ++i) a[i] += (double)h[(size_t)b * 16 + i] / grid;
    const double items = (double)total / grid, e_items = std::max(1.0, a[12]);
    fprintf(stderr, "[uptail_dbg] items/cta=%.1f cycles/item: MMA total=%.0f wait_a1_full=%.0f wait_acc1_empty=%.0f wait_acc2_empty=%.0f "
                    "wait_a2_full=%.0f | builder total=%.0f wait_a1_empty=%.0f wait_x_full=%.0f | mid-drain total=%.0f wait_a2_empty=%.0f wait_acc1_full=%.0f "
                    "tmem_ld=%.0f tmem_st=%.0f | epilogue(group 0, per own item) total=%.0f wait_acc2_full=%.0f\n",
            items, a[0] / items, a[1] / items, a[2] / items, a[3] / items, a[4] / items, a[6] / items, a[5] / items,
            a[13] / items, a[9] / items, a[7] / items, a[8] / items, a[14] / items, a[15] / items, a[11] / e_items, a[10] / e_items);
  }
  return VEX_OK;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Layout {
  size_t a, b, c, d, x5, total;  // offsets: bufA (x1/x4), bufB (x2), bufC (x3/up), bufD (extra blocks), x5
};
Layout ws_layout(const vex_cnn_desc& d, size_t npix) {
  Layout L;
  size_t off = 0;
  L.a = off; off += align_up(npix * std::max(d.c1, d.c4) * 2, 1024);
  L.b = off; off += align_up(npix * d.c2 * 2, 1024);
  L.c = off; off += align_up(npix * std::max(d.c3, 4 * d.up_ch) * 2, 1024);
  L.d = off; off += d.n_extra > 0 ? align_up(npix * d.c3 * 2, 1024) : 0;
  L.x5 = off; off += align_up(npix * 16, 1024);
  L.total = off;
  return L;
}

}  // namespace

// ------------------------------------------------------------------------------------------- weights

int cnn_load_weights(vex_ctx* ctx, const vex_cnn_desc* arch, const float* const* t, int n_tensors) {
  if (!arch || !t || arch->n_extra < 0 || arch->n_extra > VEX_CNN_MAX_EXTRA ||
      n_tensors != VEX_CNN_NUM_TENSORS + VEX_CNN_TENSORS_PER_EXTRA * arch->n_extra)
    return fail(ctx, VEX_ERR_INVALID,
                "vex_cnn_load_weights: need 30 + 6 * n_extra tensors in state_dict order, n_extra in [0,4]");
  for (int i = 0; i < n_tensors; ++i)
    if (!t[i]) return fail(ctx, VEX_ERR_INVALID, "vex_cnn_load_weights: null tensor");
  const int c1 = arch->c1, c2 = arch->c2, c3 = arch->c3, c4 = arch->c4, up = arch->up_ch;
  auto okw = [](int c) { return c >= 128 && c <= 512 && c % 128 == 0; };
  if (!okw(c1) || !okw(c2) || !okw(c3) || !okw(c4) || c1 != c2 || up != 64)
    return fail(ctx, VEX_ERR_UNSUPPORTED,
                "vex_cnn_load_weights: widths must be multiples of 128 in [128,512], c1 == c2 (skip), up_ch == 64");
  if (!ctx->cnn) ctx->cnn = new CnnPlan();
  CnnPlan* P = ctx->cnn;
  P->arch = *arch;
  if (!P->encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    VEX_CUDA(ctx, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn || q != cudaDriverEntryPointSuccess)
      return fail(ctx, VEX_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled not available in this driver");
    P->encode = fn;
  }
  // state_dict order (include/vex.h): conv1..5 {w,b}, upconv1 {w,b}, upconv2 {w,b}, bn1..4 {w,b,mean,var}
  const float *cw[5], *cb[5];
  for (int i = 0; i < 5; ++i) { cw[i] = t[2 * i]; cb[i] = t[2 * i + 1]; }
  const float *u1w = t[10], *u1b = t[11], *u2w = t[12], *u2b = t[13];
  const int cin[5] = {3, c1, c2, c3, c4}, cout[5] = {c1, c2, c3, c4, 3};
  // eval-mode BN fold (app.py:284-287, 331): y = (conv + b - mean) * g / sqrt(var + eps) + beta
  std::vector<float> scale[4], shift[4];
  for (int i = 0; i < 4; ++i) {
    const float *g = t[14 + 4 * i], *be = t[15 + 4 * i], *mu = t[16 + 4 * i], *var = t[17 + 4 * i];
    scale[i].resize(cout[i]); shift[i].resize(cout[i]);
    for (int c = 0; c < cout[i]; ++c) {
      const float s = g[c] / sqrtf(var[c] + arch->bn_eps);
      scale[i][c] = s;
      shift[i][c] = (cb[i][c] - mu[c]) * s + be[c];
    }
  }
  int rc;
  {  // conv1: [27][c1], k = (ky*3+kx)*3+ci; folds BN scale and the 1/255 of app.py:535
    std::vector<float> w((size_t)27 * c1);
    for (int co = 0; co < c1; ++co)
      for (int ci = 0; ci < 3; ++ci)
        for (int k = 0; k < 9; ++k)
          w[(size_t)(k * 3 + ci) * c1 + co] = cw[0][((size_t)co * 3 + ci) * 9 + k] * scale[0][co] * (1.0f / 255.0f);
    if ((rc = upload(ctx, &P->w1g, w.data(), w.size() * 4))) return rc;
    if ((rc = upload(ctx, &P->b1, shift[0].data(), (size_t)c1 * 4))) return rc;
  }
  for (int l = 0; l < 3; ++l) {  // conv2..conv4 on tensor cores
    TcLayer& L = P->tc[l];
    L.cin = cin[l + 1]; L.cout = cout[l + 1];
    const int want = (l == 1) ? ctx->opt_conv3_bn : 128;
    L.bn = (want == 256 && L.cout % 256 == 0) ? 256 : 128;
    L.w_folded.resize((size_t)L.cout * L.cin * 9);
    for (int co = 0; co < L.cout; ++co)
      for (size_t j = 0; j < (size_t)L.cin * 9; ++j)
        L.w_folded[(size_t)co * L.cin * 9 + j] = cw[l + 1][(size_t)co * L.cin * 9 + j] * scale[l + 1][co];
    if ((rc = pack_tc(ctx, L))) return rc;
    if ((rc = upload(ctx, &L.bias, shift[l + 1].data(), (size_t)L.cout * 4))) return rc;
  }
  for (int e = 0; e < arch->n_extra; ++e) {  // deepened plans: conv3x<e+1> c3 -> c3, BN folded like conv2..conv4
    const float* const* x = t + VEX_CNN_NUM_TENSORS + VEX_CNN_TENSORS_PER_EXTRA * e;
    const float *w = x[0], *b = x[1], *g = x[2], *be = x[3], *mu = x[4], *var = x[5];
    TcLayer& L = P->extra[e];
    L.cin = c3; L.cout = c3;
    L.bn = (ctx->opt_conv3_bn == 256 && c3 % 256 == 0) ? 256 : 128;
    L.w_folded.resize((size_t)c3 * c3 * 9);
    std::vector<float> sh(c3);
    for (int co = 0; co < c3; ++co) {
      const float sc = g[co] / sqrtf(var[co] + arch->bn_eps);
      sh[co] = (b[co] - mu[co]) * sc + be[co];
      for (size_t j = 0; j < (size_t)c3 * 9; ++j)
        L.w_folded[(size_t)co * c3 * 9 + j] = w[(size_t)co * c3 * 9 + j] * sc;
    }
    if ((rc = pack_tc(ctx, L))) return rc;
    if ((rc = upload(ctx, &L.bias, sh.data(), (size_t)c3 * 4))) return rc;
  }
  {  // conv5: [c4/64][9][64][4]
    std::vector<float> w((size_t)c4 * 9 * 4, 0.f);
    for (int co = 0; co < 3; ++co)
      for (int ci = 0; ci < c4; ++ci)
        for (int k = 0; k < 9; ++k)
          w[(((size_t)(ci / 64) * 9 + k) * 64 + (ci % 64)) * 4 + co] = cw[4][((size_t)co * c4 + ci) * 9 + k];
    if ((rc = upload(ctx, &P->w5g, w.data(), w.size() * 4))) return rc;
    if ((rc = upload(ctx, &P->b5, cb[4], 12))) return rc;
  }
  {  // upconv1 (ConvTranspose2d weight [ci=3][co=up][4][4]) -> [phase][(a*2+b)*3+ci][co]
     // (ky, dy) per phase bit p and tap index a:  p=0: {(1,0), (3,-1)}   p=1: {(0,+1), (2,0)}
    const int kys[2][2] = {{1, 3}, {0, 2}};
    std::vector<float> w((size_t)48 * up);
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px)
        for (int a = 0; a < 2; ++a)
          for (int b = 0; b < 2; ++b)
            for (int ci = 0; ci < 3; ++ci)
              for (int co = 0; co < up; ++co)
                w[((size_t)(py * 2 + px) * 12 + (a * 2 + b) * 3 + ci) * up + co] =
                    u1w[(((size_t)ci * up + co) * 4 + kys[py][a]) * 4 + kys[px][b]];
    if ((rc = upload(ctx, &P->wu1, w.data(), w.size() * 4))) return rc;
    if ((rc = upload(ctx, &P->bu1, u1b, (size_t)up * 4))) return rc;
  }
  {  // upconv2 (ConvTranspose2d weight [ci=up][co=3][3][3], s1 p1) == conv3x3 with the kernel flipped:
     // tap (ky,kx) of the equivalent conv uses w[ci][co][2-ky][2-kx]
    std::vector<float> w((size_t)up * 9 * 4, 0.f);
    for (int co = 0; co < 3; ++co)
      for (int ci = 0; ci < up; ++ci)
        for (int ky = 0; ky < 3; ++ky)
          for (int kx = 0; kx < 3; ++kx)
            w[(((size_t)(ci / 64) * 9 + ky * 3 + kx) * 64 + (ci % 64)) * 4 + co] =
                u2w[(((size_t)ci * 3 + co) * 3 + (2 - ky)) * 3 + (2 - kx)];
    if ((rc = upload(ctx, &P->wu2, w.data(), w.size() * 4))) return rc;
    if ((rc = upload(ctx, &P->bu2, u2b, 12))) return rc;
  }
  // ---- tensor-core images of the thin layers (thin_tc.cuh); k = (ky*3+kx)*3+ci, rows = output columns
  if (c1 <= 256) {   // conv1: [c1 rows][k < 27], BN scale and 1/255 folded
    rc = pack_rows_sw128(ctx, &P->t1w, c1, 1, [&](int r, int k) -> float {
      if (k == 27 || k == 28) {                      // folded bias as hi + lo bf16 against the builders' constant-1 columns
        const uint32_t hb = (uint32_t)f32_to_bf16_rne(shift[0][r]) << 16;
        float hi;
        memcpy(&hi, &hb, 4);
        return k == 27 ? hi : shift[0][r] - hi;
      }
      if (k > 28) return 0.f;
      const int ci = k % 3, tap = k / 3;
      return cw[0][((size_t)r * 3 + ci) * 9 + tap] * scale[0][r] * (1.0f / 255.0f);
    });
    if (rc) return rc;
  }
  if (c1 <= 256) {   // conv1, fp16 offset form (thin_tc.cuh MODE_CONV1_H): k = 10 ky + 3 kx + cb with cb the BUFFER channel
    for (int bgr = 0; bgr < 2; ++bgr) {
      std::vector<uint16_t> img((size_t)c1 * 64, 0);
      for (int r = 0; r < c1; ++r) {
        double corr = 0.0;
        uint16_t row[32] = {0};
        for (int ky = 0; ky < 3; ++ky)
          for (int j = 0; j < 9; ++j) {
            const int kx = j / 3, cb = j % 3, ci = bgr ? 2 - cb : cb;
            const float wv = cw[0][((size_t)r * 3 + ci) * 9 + ky * 3 + kx] * scale[0][r] * (1.0f / 255.0f);
            const __half hw = __float2half_rn(wv);
            memcpy(&row[10 * ky + j], &hw, 2);
            corr += 1024.0 * (double)__half2float(hw);                    // what the operand offset adds to the sum
          }
        const float bp = (float)((double)shift[0][r] - corr);
        const __half bh = __float2half_rn(bp), bl = __float2half_rn(bp - __half2float(bh));
        memcpy(&row[30], &bh, 2);
        memcpy(&row[31], &bl, 2);
        for (int k = 0; k < 32; ++k) img[(size_t)r * 64 + (((k >> 3) ^ (r & 7)) << 3) + (k & 7)] = row[k];
      }
      if ((rc = upload(ctx, &P->t1wh[bgr], img.data(), img.size() * 2))) return rc;
    }
  }
  if (c1 <= 256) {   // conv1 for fp32 input: the input value arrives as hi (k < 27) + lo (k = 32..58) bf16, same weight for both
    rc = pack_rows_sw128(ctx, &P->t1wf, c1, 1, [&](int r, int k) -> float {
      if (k == 27 || k == 28) {
        const uint32_t hb = (uint32_t)f32_to_bf16_rne(shift[0][r]) << 16;
        float hi;
        memcpy(&hi, &hb, 4);
        return k == 27 ? hi : shift[0][r] - hi;
      }
      const int kk = k >= 32 ? k - 32 : k;
      if (kk >= 27) return 0.f;
      const int ci = kk % 3, tap = kk / 3;
      return cw[0][((size_t)r * 3 + ci) * 9 + tap] * scale[0][r];
    });
    if (rc) return rc;
  }
  {  // upconv1 as conv3x3 3 -> 4*up + PixelShuffle(2): row = phase*up + co, phase = py*2+px.
     // 3x3 tap index ky' (input row y + ky' - 1) -> transposed-conv kernel row ky:  py=0: {1->1, 0->3}, py=1: {2->0, 1->2}
    const int kmap[2][3] = {{3, 1, -1}, {-1, 2, 0}};
    // the bias is folded in as two bf16 terms (hi + lo) at k = 27, 28; the builders put 1.0 there (thin_tc.cuh, tail_tc.cuh)
    rc = pack_rows_sw128(ctx, &P->tu1wb, 4 * up, 1, [&](int r, int k) -> float {
      const int ph = r / up, co = r % up;
      if (k == 27 || k == 28) {
        const uint32_t hb = (uint32_t)f32_to_bf16_rne(u1b[co]) << 16;
        float hi;
        memcpy(&hi, &hb, 4);
        return k == 27 ? hi : u1b[co] - hi;
      }
      if (k > 28) return 0.f;
      const int ci = k % 3, tap = k / 3;
      const int ky = kmap[ph >> 1][tap / 3], kx = kmap[ph & 1][tap % 3];
      if (ky < 0 || kx < 0) return 0.f;
      return u1w[(((size_t)ci * up + co) * 4 + ky) * 4 + kx];
    });
    if (rc) return rc;
  }
  // conv5 / upconv2 as 1x1 GEMMs: row n = tap*3+co (27 of 32), K = input channels
  rc = pack_rows_sw128(ctx, &P->t5w, 32, c4 / 64, [&](int r, int k) -> float {
    if (r >= 27) return 0.f;
    return cw[4][((size_t)(r % 3) * c4 + k) * 9 + r / 3];
  });
  if (rc) return rc;
  rc = pack_rows_sw128(ctx, &P->tu2w, 32, up / 64, [&](int r, int k) -> float {
    if (r >= 27) return 0.f;
    const int co = r % 3, tap = r / 3, ky = tap / 3, kx = tap % 3;
    return u2w[(((size_t)k * 3 + co) * 3 + (2 - ky)) * 3 + (2 - kx)];
  });
  if (rc) return rc;
  if (up == 64) {   // tail_tc3.cuh: slice dy (the "chunk" of pack_rows_sw128), row dx*3 + co, k = up channel
    rc = pack_rows_sw128(ctx, &P->tu2w3, 16, 3, [&](int r, int kq) -> float {
      if (r >= 9) return 0.f;
      const int dy = kq / 64, k = kq % 64, dx = r / 3, co = r % 3;
      return u2w[(((size_t)k * 3 + co) * 3 + (2 - dy)) * 3 + (2 - dx)];
    });
    if (rc) return rc;
  }
  return VEX_OK;
}

// ------------------------------------------------------------------------------------------- forward

namespace {
int fold_forward(vex_ctx* ctx, const std::vector<cudaEvent_t>& ev);
}

size_t cnn_workspace_bytes(const vex_ctx* ctx, int n, int h, int w) {
  if (!ctx->cnn) return 0;
  return ws_layout(ctx->cnn->arch, (size_t)n * h * w).total;
}

static int forward_chunk(vex_ctx* ctx, const uint8_t* in, uint8_t* out_u8, float* out_f32, int tap, int n, int H,
                         int W, int in_bgr, int out_bgr, cudaStream_t s) {
  CnnPlan* P = ctx->cnn;
  const vex_cnn_desc& d = P->arch;
  const size_t npix = (size_t)n * H * W;
  const Layout L = ws_layout(d, npix);
  if (L.total > P->ws_bytes) {
    VEX_CUDA(ctx, cudaDeviceSynchronize());
    if (P->ws) cudaFree(P->ws);
    P->ws = nullptr; P->ws_bytes = 0;
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&P->ws), L.total));
    P->ws_bytes = L.total;
  }
  __nv_bfloat16* x1 = reinterpret_cast<__nv_bfloat16*>(P->ws + L.a);
  __nv_bfloat16* x2 = reinterpret_cast<__nv_bfloat16*>(P->ws + L.b);
  __nv_bfloat16* x3 = reinterpret_cast<__nv_bfloat16*>(P->ws + L.c);
  __nv_bfloat16* x4 = x1;
  __nv_bfloat16* up = x3;                     // bufC; free again once conv4 has run
  float* x5 = reinterpret_cast<float*>(P->ws + L.x5);
  const int HW = H * W;
  auto tap_bf16 = [&](const __nv_bfloat16* x, int C) -> int {
    const size_t items = npix * (C / 8);
    small::nhwc_bf16_to_nchw_f32_kernel<<<(unsigned)((items + 255) / 256), 256, 0, s>>>(x, out_f32, npix, HW, C);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    return VEX_OK;
  };
  auto mark = [&]() {                      // layer boundary for vex_cnn_profile_ms
    if (!ctx->opt_profile) return;
    cudaEvent_t e;
    if (!P->prof_pool.empty()) { e = P->prof_pool.back(); P->prof_pool.pop_back(); }
    else if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, s);
    P->prof.push_back(e);
  };
  // option "nvtx": one range per layer launch (tracing); ranges close on every exit path
  struct NvtxLayers {
    bool on, open = false;
    explicit NvtxLayers(bool e) : on(e) {}
    void operator()(const char* name) { if (!on) return; if (open) nvtxRangePop(); nvtxRangePushA(name); open = true; }
    ~NvtxLayers() { if (open) nvtxRangePop(); }
  } layer(ctx->opt_nvtx != 0);
  int rc;
  const bool thin_tc = ctx->opt_thin_tc != 0;
  mark();
  layer("vex.conv1");
  if (in_bgr == IN_F32_NCHW) {
    if (d.c1 == 128) rc = launch_i27<thin::MODE_CONV1_F32, 128>(ctx, 0x800u, in, P->t1wf, x1, n, H, W, 0, s);
    else if (d.c1 == 256) rc = launch_i27<thin::MODE_CONV1_F32, 256>(ctx, 0x1000u, in, P->t1wf, x1, n, H, W, 0, s);
    else rc = fail(ctx, VEX_ERR_UNSUPPORTED, "vex_cnn2x_f32in: c1 must be 128 or 256");
    if (rc) return rc;
  } else if (thin_tc && ctx->opt_conv1_variant == 0 && d.c1 == 128) {
    if ((rc = launch_i27<thin::MODE_CONV1_H, 128>(ctx, 0x2000u, in, P->t1wh[in_bgr ? 1 : 0], x1, n, H, W, 0, s))) return rc;
  } else if (thin_tc && ctx->opt_conv1_variant == 0 && d.c1 == 256) {
    if ((rc = launch_i27<thin::MODE_CONV1_H, 256>(ctx, 0x4000u, in, P->t1wh[in_bgr ? 1 : 0], x1, n, H, W, 0, s))) return rc;
  } else if (thin_tc && d.c1 == 128) {
    if ((rc = launch_i27<thin::MODE_CONV1, 128>(ctx, 8u, in, P->t1w, x1, n, H, W, in_bgr, s))) return rc;
  } else if (thin_tc && d.c1 == 256) {
    if ((rc = launch_i27<thin::MODE_CONV1, 256>(ctx, 16u, in, P->t1w, x1, n, H, W, in_bgr, s))) return rc;
  } else {
    dim3 grid((W + 31) / 32, (H + 7) / 8, n);
    small::conv1_u8_kernel<<<grid, dim3(32, 8), small::conv1_smem(d.c1), s>>>(in, x1, P->w1g, P->b1, H, W, d.c1,
                                                                              in_bgr);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
  }
  mark();
  if (out_f32 && tap == 1) return tap_bf16(x1, d.c1);
  layer("vex.conv2");
  if ((rc = run_tc(ctx, P->tc[0], x1, x2, x1, n, H, W, s))) return rc;        // + x1 after ReLU (app.py:292)
  mark();
  if (out_f32 && tap == 2) return tap_bf16(x2, d.c2);
  layer("vex.conv3");
  if ((rc = run_tc(ctx, P->tc[1], x2, x3, nullptr, n, H, W, s))) return rc;
  if (out_f32 && tap == 3) return tap_bf16(x3, d.c3);
  for (int e = 0; e < d.n_extra; ++e) {                                        // ping-pong bufC <-> bufD
    __nv_bfloat16* nxt = reinterpret_cast<__nv_bfloat16*>(P->ws + ((e & 1) ? L.c : L.d));
    if ((rc = run_tc(ctx, P->extra[e], x3, nxt, nullptr, n, H, W, s))) return rc;
    x3 = nxt;
    if (out_f32 && tap == 7 + e) return tap_bf16(x3, d.c3);
  }
  mark();                                                                      // conv3's slot includes the extra blocks
  layer("vex.conv4");
  if ((rc = run_tc(ctx, P->tc[2], x3, x4, nullptr, n, H, W, s))) return rc;
  mark();
  if (out_f32 && tap == 4) return tap_bf16(x4, d.c4);
  if (!(ctx->attr_mask & 4u)) {
    VEX_CUDA(ctx, cudaFuncSetAttribute(small::conv3x3_to3_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)small::T3_SMEM));
    VEX_CUDA(ctx, cudaFuncSetAttribute(small::conv3x3_to3_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)small::T3_SMEM));
    ctx->attr_mask |= 4u;
  }
  layer("vex.conv5");
  if (thin_tc) {
    if ((rc = launch_g27<thin::MODE_CONV5>(ctx, 32u, x4, d.c4, P->t5w, P->b5, x5, nullptr, nullptr, n, H, W, 0, s)))
      return rc;
  } else {
    dim3 grid((W + 31) / 32, (H + small::T3_TILE_H - 1) / small::T3_TILE_H, n);
    small::conv3x3_to3_kernel<0><<<grid, dim3(32, 8), small::T3_SMEM, s>>>(x4, P->w5g, P->b5, x5, nullptr, nullptr,
                                                                           H, W, d.c4, 0);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
  }
  mark();
  if (out_f32 && tap == 5) {
    small::nhw4_f32_to_nchw_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, s>>>(x5, out_f32, npix, HW);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    return VEX_OK;
  }
  layer("vex.tail");
  if (thin_tc && ctx->opt_fuse_tail && d.up_ch == 64 && !(out_f32 && tap == 6)) {
    // upconv1 + relu + upconv2 + clamp + uint8 in one kernel: `up` never reaches HBM
    if ((rc = launch_uptail(ctx, x5, P->tu1wb, ctx->opt_fuse_tail == 4 ? P->tu2w3 : P->tu2w, P->bu2, out_u8, out_f32, n, H, W,
                            out_bgr, s)))
      return rc;
    mark();                                  // the fused time is reported in upconv1's slot, upconv2's reads 0
    mark();
    return VEX_OK;
  }
  if (thin_tc) {
    if ((rc = launch_i27<thin::MODE_UP1, 256>(ctx, 64u, x5, P->tu1wb, up, n, H, W, 0, s))) return rc;
  } else {
    dim3 grid((W + 31) / 32, H, n);
    small::upconv1_kernel<<<grid, 128, (size_t)49 * d.up_ch * 4, s>>>(x5, up, P->wu1, P->bu1, H, W, d.up_ch);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
  }
  mark();
  if (out_f32 && tap == 6) {
    const size_t items = npix * 4 * (d.up_ch / 8);
    small::nhwc_bf16_to_nchw_f32_kernel<<<(unsigned)((items + 255) / 256), 256, 0, s>>>(up, out_f32, npix * 4,
                                                                                         4 * HW, d.up_ch);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    return VEX_OK;
  }
  if (thin_tc) {
    if ((rc = launch_g27<thin::MODE_UP2>(ctx, 128u, up, d.up_ch, P->tu2w, P->bu2, nullptr, out_u8, out_f32, n, 2 * H,
                                         2 * W, out_bgr, s)))
      return rc;
  } else {
    const int H2 = 2 * H, W2 = 2 * W;
    dim3 grid((W2 + 31) / 32, (H2 + small::T3_TILE_H - 1) / small::T3_TILE_H, n);
    small::conv3x3_to3_kernel<1><<<grid, dim3(32, 8), small::T3_SMEM, s>>>(up, P->wu2, P->bu2, nullptr, out_u8,
                                                                           out_f32, H2, W2, d.up_ch, out_bgr);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
  }
  mark();
  return VEX_OK;
}

int cnn_forward(vex_ctx* ctx, const uint8_t* in, uint8_t* out_u8, float* out_f32, int tap, int n, int h, int w,
                int in_bgr, int out_bgr, cudaStream_t s) {
  if (!ctx->cnn || !ctx->cnn->w1g)
    return fail(ctx, VEX_ERR_NO_WEIGHTS, "vex_cnn_load_weights has not been called");
  if ((long long)h * w > (1LL << 28)) return fail(ctx, VEX_ERR_INVALID, "frame too large");
  // bound the activation workspace: sub-batches of at most ws_budget bytes (at least one frame)
  {
    CnnPlan* P = ctx->cnn;
    if (!P->prof.empty()) {                   // keep the previous forward's events until they are folded into the statistics
      if (P->prof_folded || !ctx->opt_profile) {
        for (cudaEvent_t e : P->prof) P->prof_pool.push_back(e);
      } else {
        P->prof_done.push_back(P->prof);
        if (P->prof_done.size() > 512) {      // bound the event count: fold the oldest (long finished) forward
          int rc = fold_forward(ctx, P->prof_done.front());
          if (rc) return rc;
          for (cudaEvent_t e : P->prof_done.front()) P->prof_pool.push_back(e);
          P->prof_done.erase(P->prof_done.begin());
        }
      }
      P->prof.clear();
    }
    P->prof_folded = false;
  }
  // One activation workspace per ctx: a forward on stream s must not start before the previous forward, on whatever
  // stream that ran (the caller's for vex_cnn2x_u8, the ctx's compute stream for the host-buffer entries), has ended.
  if (!ctx->ev_ws) VEX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_ws, cudaEventDisableTiming));
  if (ctx->ws_used) VEX_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_ws, 0));
  struct WsMark {                             // recorded on every exit path, after the last launch
    vex_ctx* c; cudaStream_t st;
    ~WsMark() { if (cudaEventRecord(c->ev_ws, st) == cudaSuccess) c->ws_used = true; }
  } ws_mark{ctx, s};
  const size_t per_frame = ws_layout(ctx->cnn->arch, (size_t)h * w).total;
  const size_t budget = (size_t)ctx->opt_ws_budget_mb << 20;
  int sub = (int)std::max<size_t>(1, std::min<size_t>((size_t)n, budget / std::max<size_t>(per_frame, 1)));
  const vex_cnn_desc& d = ctx->cnn->arch;
  if (tap > 6 + d.n_extra) return fail(ctx, VEX_ERR_INVALID, "vex_cnn2x_f32_debug: tap beyond the last extra block");
  const size_t out_ch[7 + VEX_CNN_MAX_EXTRA] = {3, (size_t)d.c1, (size_t)d.c2, (size_t)d.c3, (size_t)d.c4, 3, (size_t)d.up_ch,
                                                (size_t)d.c3, (size_t)d.c3, (size_t)d.c3, (size_t)d.c3};
  const size_t out_px = (tap == 0 || tap == 6) ? (size_t)4 * h * w : (size_t)h * w;
  for (int f0 = 0; f0 < n; f0 += sub) {
    const int nf = std::min(sub, n - f0);
    int rc = forward_chunk(ctx, in + (size_t)f0 * h * w * 3 * (in_bgr == IN_F32_NCHW ? 4 : 1), out_u8 ? out_u8 + (size_t)f0 * h * w * 12 : nullptr,
                           out_f32 ? out_f32 + (size_t)f0 * out_ch[tap] * out_px : nullptr, tap, nf, h, w, in_bgr,
                           out_bgr, s);
    if (rc) return rc;
  }
  return VEX_OK;
}

// ------------------------------------------------------------------------------------------- no-allocation entry

// Set-up half (called from vex_cnn_load_weights with the ctx's device current): resolve this device's copy of the
// static scratch.  A failure is kept and reported by the entry point; weight loading itself is not affected.
void cnn_noalloc_setup(vex_ctx* ctx) {
  CnnPlan* P = ctx->cnn;
  if (!P || P->noalloc_ws) return;
  void* ptr = nullptr;
  P->noalloc_err = cudaGetSymbolAddress(&ptr, g_noalloc_ws);
  if (P->noalloc_err != cudaSuccess) { cudaGetLastError(); ptr = nullptr; }
  P->noalloc_ws = static_cast<uint8_t*>(ptr);
}

// The forward of cnn_forward with the static scratch as the activation workspace: same kernels, same launch order, so the
// output is the existing entry point's bit for bit.  No cudaMalloc / cudaFree / event creation on this path: the options
// that would allocate (profile events, conv_dbg counters) are refused.  Every layer writes its output before the next
// one reads it and image borders come from the tensor maps' out-of-bounds fill, so nothing in the scratch has to start
// at zero and the previous call's contents are harmless.  All calls on a device share the one array: calls that overlap
// on the GPU (different streams without an ordering between them, or two contexts of one process) race.
int cnn_forward_noalloc(vex_ctx* ctx, const uint8_t* in, uint8_t* out_u8, int n, int h, int w, int in_bgr, int out_bgr,
                        cudaStream_t s) {
  CnnPlan* P = ctx->cnn;
  if (!P || !P->w1g) return fail(ctx, VEX_ERR_NO_WEIGHTS, "vex_cnn_load_weights has not been called");
  if (!P->noalloc_ws)
    return fail(ctx, VEX_ERR_CUDA, std::string("vex_cnn2x_u8_noalloc: cudaGetSymbolAddress of the static scratch failed: ") +
                                       cudaGetErrorString(P->noalloc_err));
  if (ctx->opt_profile || ctx->opt_conv_dbg)
    return fail(ctx, VEX_ERR_UNSUPPORTED, "vex_cnn2x_u8_noalloc: options profile / conv_dbg allocate; switch them off");
  if ((long long)h * w > (1LL << 28)) return fail(ctx, VEX_ERR_INVALID, "frame too large");
  const size_t per_frame = ws_layout(P->arch, (size_t)h * w).total;
  if (per_frame > (size_t)VEX_NOALLOC_SCRATCH_BYTES)
    return fail(ctx, VEX_ERR_NOMEM, "vex_cnn2x_u8_noalloc: a " + std::to_string(h) + " x " + std::to_string(w) + " frame needs " +
                                        std::to_string(per_frame) + " bytes of scratch, the built-in scratch holds " +
                                        std::to_string((size_t)VEX_NOALLOC_SCRATCH_BYTES) +
                                        " (VEX_NOALLOC_SCRATCH_BYTES); use vex_cnn2x_u8");
  // frames that fit the scratch together go as one sub-batch; forward_chunk finds ws_bytes large enough and allocates nothing
  int sub = (int)std::min<size_t>((size_t)n, (size_t)VEX_NOALLOC_SCRATCH_BYTES / per_frame);
  while (sub > 1 && ws_layout(P->arch, (size_t)sub * h * w).total > (size_t)VEX_NOALLOC_SCRATCH_BYTES) --sub;
  struct Swap {                               // the ctx's own workspace comes back on every exit path
    CnnPlan* p; uint8_t* ws; size_t bytes;
    ~Swap() { p->ws = ws; p->ws_bytes = bytes; }
  } swap{P, P->ws, P->ws_bytes};
  P->ws = P->noalloc_ws;
  P->ws_bytes = (size_t)VEX_NOALLOC_SCRATCH_BYTES;
  for (int f0 = 0; f0 < n; f0 += sub) {
    const int nf = std::min(sub, n - f0);
    const int rc = forward_chunk(ctx, in + (size_t)f0 * h * w * 3, out_u8 + (size_t)f0 * h * w * 12, nullptr, 0, nf, h, w,
                                 in_bgr, out_bgr, s);
    if (rc) return rc;
  }
  return VEX_OK;
}

namespace {
// per-layer ms of one forward (8 events per sub-batch, sub-batches summed); synchronises on its last event
int forward_ms(vex_ctx* ctx, const std::vector<cudaEvent_t>& ev, float* ms7) {
  VEX_CUDA(ctx, cudaEventSynchronize(ev.back()));
  for (int i = 0; i < 7; ++i) ms7[i] = 0.f;
  for (size_t b = 0; b + 8 <= ev.size(); b += 8)
    for (int i = 0; i < 7; ++i) {
      float ms = 0.f;
      VEX_CUDA(ctx, cudaEventElapsedTime(&ms, ev[b + i], ev[b + i + 1]));
      ms7[i] += ms;
    }
  return VEX_OK;
}
int fold_forward(vex_ctx* ctx, const std::vector<cudaEvent_t>& ev) {
  CnnPlan* P = ctx->cnn;
  if (ev.empty() || ev.size() % 8 != 0) return VEX_OK;     // a debug-tap forward that returned early: not a full forward
  float ms[7];
  int rc = forward_ms(ctx, ev, ms);
  if (rc) return rc;
  for (int i = 0; i < 7; ++i) {
    P->prof_sum[i] += ms[i];
    P->prof_min[i] = P->prof_n ? std::min(P->prof_min[i], ms[i]) : ms[i];
    P->prof_max[i] = P->prof_n ? std::max(P->prof_max[i], ms[i]) : ms[i];
  }
  P->prof_n++;
  return VEX_OK;
}
}  // namespace

int cnn_profile_ms(vex_ctx* ctx, float* ms7) {
  CnnPlan* P = ctx->cnn;
  if (!P || P->prof.empty() || P->prof.size() % 8 != 0)
    return fail(ctx, VEX_ERR_INVALID, "vex_cnn_profile_ms: no complete profiled forward (set option profile=1)");
  return forward_ms(ctx, P->prof, ms7);
}

void cnn_profile_reset(vex_ctx* ctx) {
  CnnPlan* P = ctx->cnn;
  if (!P) return;
  for (auto& g : P->prof_done)
    for (cudaEvent_t e : g) P->prof_pool.push_back(e);
  P->prof_done.clear();
  for (cudaEvent_t e : P->prof) P->prof_pool.push_back(e);
  P->prof.clear();
  P->prof_folded = false;
  P->prof_n = 0;
  for (int i = 0; i < 7; ++i) { P->prof_sum[i] = 0; P->prof_min[i] = P->prof_max[i] = 0.f; }
}

int cnn_profile_stats(vex_ctx* ctx, float* mean7, float* min7, float* max7, int* n_forwards) {
  CnnPlan* P = ctx->cnn;
  if (!P) return fail(ctx, VEX_ERR_INVALID, "vex_cnn_profile_stats: no CNN plan");
  for (auto& g : P->prof_done) {
    int rc = fold_forward(ctx, g);
    if (rc) return rc;
    for (cudaEvent_t e : g) P->prof_pool.push_back(e);
  }
  P->prof_done.clear();
  if (!P->prof_folded && !P->prof.empty()) {
    int rc = fold_forward(ctx, P->prof);
    if (rc) return rc;
    P->prof_folded = true;
  }
  *n_forwards = P->prof_n;
  for (int i = 0; i < 7; ++i) {
    if (mean7) mean7[i] = P->prof_n ? (float)(P->prof_sum[i] / P->prof_n) : 0.f;
    if (min7) min7[i] = P->prof_min[i];
    if (max7) max7[i] = P->prof_max[i];
  }
  return VEX_OK;
}

int cnn_repack(vex_ctx* ctx) {
  CnnPlan* P = ctx->cnn;
  if (!P || !P->w1g) return VEX_OK;
  TcLayer* layers[1 + VEX_CNN_MAX_EXTRA] = {&P->tc[1], &P->extra[0], &P->extra[1], &P->extra[2], &P->extra[3]};
  for (int i = 0; i < 1 + P->arch.n_extra; ++i) {
    TcLayer& L = *layers[i];
    const int bn = (ctx->opt_conv3_bn == 256 && L.cout % 256 == 0) ? 256 : 128;
    if (bn == L.bn) continue;
    VEX_CUDA(ctx, cudaDeviceSynchronize());
    L.bn = bn;
    int rc = pack_tc(ctx, L);
    if (rc) return rc;
  }
  return VEX_OK;
}

void cnn_destroy(vex_ctx* ctx) {
  CnnPlan* P = ctx->cnn;
  if (!P) return;
  float* fs[] = {P->w1g, P->b1, P->w5g, P->b5, P->wu1, P->bu1, P->wu2, P->bu2};
  uint8_t* us[] = {P->t1w, P->tu1wb, P->t5w, P->tu2w, P->tu2w3, P->t1wf, P->t1wh[0], P->t1wh[1]};
  for (uint8_t* u : us)
    if (u) cudaFree(u);
  for (float* f : fs)
    if (f) cudaFree(f);
  for (TcLayer& L : P->tc) {
    if (L.wpack) cudaFree(L.wpack);
    if (L.bias) cudaFree(L.bias);
  }
  for (TcLayer& L : P->extra) {
    if (L.wpack) cudaFree(L.wpack);
    if (L.bias) cudaFree(L.bias);
  }
  if (P->ws) cudaFree(P->ws);
  for (cudaEvent_t e : P->prof) cudaEventDestroy(e);
  for (auto& g : P->prof_done)
    for (cudaEvent_t e : g) cudaEventDestroy(e);
  for (cudaEvent_t e : P->prof_pool) cudaEventDestroy(e);
  delete P;
  ctx->cnn = nullptr;
}

}  // namespace vex
