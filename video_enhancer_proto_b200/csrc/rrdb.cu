// Real-ESRGAN RRDBNet (reference models/enhancement/vsr/realesrgan_handler.py:23-100, frame path :286-372) on sm_100a:
// weight packing, workspace, the chain of ~350 layer launches per image, and the reference's tile loop on the device.
//
//   u8 HWC --conv_first--> feat (fp16 NHWC 64) --copy--> D0[:, 0:64]
//   per RRDBBlock (dense buffers A, B, C of 192 channels, rotating):
//        rdb1 on A: conv1..4 -> A[64:192] (LeakyReLU), conv5 -> B[0:64] = acc*0.2 + A[0:64]
//        rdb2 on B:                                     conv5 -> C[0:64] = acc*0.2 + B[0:64]
//        rdb3 on C:                                     conv5 -> B[0:64] = (acc*0.2 + C[0:64])*0.2 + A[0:64]     (A,B,C) <- (B,C,A)
//   conv_body: trunk = acc + feat, stored 2x2-replicated (= F.interpolate nearest of :93) -> U1 [2H][2W][64]
//   conv_up1 + LeakyReLU (scale 4: stored 2x2-replicated again -> [4H][4W]) ; conv_up2 + LeakyReLU (scale 4 only)
//   conv_hr + LeakyReLU ; conv_last -> clamp(0,1) -> x255 -> truncate -> u8 HWC (sH, sW)
#include <algorithm>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>

#include <cuda_fp16.h>

#include "rrdb_tc.cuh"
#include "thin_tc.cuh"
#include "vex_internal.cuh"

namespace vex {

struct RConv {
  int cin = 0, cout = 0, nc = 0;          // nc = 64-channel chunks read (cin rounded up)
  uint8_t* wpack = nullptr;
  float* bias = nullptr;
};

struct RrdbPlan {
  vex_rrdb_desc arch{};
  uint8_t* first_w = nullptr;             // conv_first: im2col27 image (64 rows x 128 B, bf16 hi + lo)
  std::vector<RConv> rdb;                 // num_block * 3 * 5
  RConv body, up1, up2, hr;
  uint8_t* last_w = nullptr;              // conv_last: gemm27 image (32 rows x 128 B, fp16)
  float* last_b = nullptr;
  uint8_t* ws = nullptr;
  size_t ws_bytes = 0;
  void* encode = nullptr;
  unsigned attr_mask = 0;
  // tile loop scratch (vex_rrdb_frame_u8)
  uint8_t* tile_in = nullptr;
  uint8_t* tile_out = nullptr;
  size_t tile_in_bytes = 0, tile_out_bytes = 0;
  // canvas batching: valid (h, w) of every image of the forward that is being enqueued (empty = all images fill the canvas)
  std::vector<int> canvas_dims;
  unsigned launch_parity = 0;             // alternates the sweep direction of consecutive conv launches (Params::reverse)
  int canvas_h0 = 0;                      // the canvas height at the network's input resolution (layers at 2x / 4x scale the table)
};

namespace {

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

uint16_t f32_to_f16(float f) {
  const __half h = __float2half_rn(f);
  uint16_t u;
  memcpy(&u, &h, 2);
  return u;
}
uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
float bf16_to_f32(uint16_t h) {
  const uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

template <typename T>
int upload(vex_ctx* ctx, T** dst, const void* src, size_t bytes) {
  if (*dst) cudaFree(*dst);
  *dst = nullptr;
  VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(dst), bytes));
  VEX_CUDA(ctx, cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice));
  return VEX_OK;
}

// nn.Conv2d weight [cout][cin][3][3] -> per block of 32 output channels [chunk][dy] slices of 96 rows x 64 fp16 in the
// SWIZZLE_128B K-major image (16-byte chunk j of row r at j ^ (r & 7)): row dx*32 + co holds W[co][ci][dy][dx], the three
// horizontal taps side by side in the MMA's N (rrdb_tc.cuh); input channels beyond cin are zero columns.
int pack_rconv(vex_ctx* ctx, RConv& L, const float* w, const float* b) {
  L.nc = (L.cin + 63) / 64;
  const int nblk = L.cout / rrdb::CO;
  std::vector<uint16_t> img((size_t)nblk * L.nc * 3 * rrdb::NB * 64, 0);
  for (int blk = 0; blk < nblk; ++blk)
    for (int c = 0; c < L.nc; ++c)
      for (int dy = 0; dy < 3; ++dy) {
        uint16_t* slice = img.data() + (((size_t)blk * L.nc + c) * 3 + dy) * rrdb::NB * 64;
        for (int r = 0; r < rrdb::NB; ++r)
          for (int k = 0; k < 64; ++k) {
            const int ci = c * 64 + k, dx = r / rrdb::CO, co = blk * rrdb::CO + r % rrdb::CO;
            const float v = ci < L.cin ? w[((size_t)co * L.cin + ci) * 9 + dy * 3 + dx] : 0.f;
            slice[(size_t)r * 64 + (((k >> 3) ^ (r & 7)) << 3) + (k & 7)] = f32_to_f16(v);
          }
      }
  int rc = upload(ctx, &L.wpack, img.data(), img.size() * 2);
  if (rc) return rc;
  return upload(ctx, &L.bias, b, (size_t)L.cout * 4);
}

// promo128: L2 promotion of 128 instead of 256 bytes.  A layer that reads only the first 64 or 128 channels of a
// 192-channel (384-byte) pixel would otherwise pull the neighbouring 128-byte chunk it never uses into L2 with every
// request: measured with ncu on the batched 1080p pass, conv1 read 439 MB for 287 MB of input, conv2/3 732 for 574.
int make_map(vex_ctx* ctx, CUtensorMap* map, const void* base, int n, int H, int W, int cs, int box_rows,
             bool promo128 = false, bool half = false) {
  RrdbPlan* P = ctx->rrdb;
  const cuuint64_t dims[4] = {(cuuint64_t)cs, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
  const cuuint64_t strides[3] = {(cuuint64_t)cs * 2, (cuuint64_t)W * cs * 2, (cuuint64_t)H * W * cs * 2};
  const cuuint32_t box[4] = {half ? 32u : 64u, 32, (cuuint32_t)box_rows, 1};   // half: 32 channels, 64-byte rows
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = reinterpret_cast<EncodeFn>(P->encode)(
      map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
      CU_TENSOR_MAP_INTERLEAVE_NONE, half ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
      promo128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(ctx, VEX_ERR_CUDA, "cuTensorMapEncodeTiled (rrdb) failed (" + std::to_string((int)r) + ")");
  return VEX_OK;
}

struct Epi {
  __half* out; int out_cs, out_c0;
  const __half* r1 = nullptr; int r1_cs = 0; float a1 = 1.f;
  const __half* r2 = nullptr; int r2_cs = 0; float a2 = 1.f;
  int lrelu = 0, ups2 = 0;
};

// one launch per layer; a 64-channel layer carries its two 32-channel halves as interleaved work items (Params::nblk)
int launch_rconv(vex_ctx* ctx, const RConv& L, const __half* in, int in_cs, const Epi& e, int n, int H, int W,
                 cudaStream_t s) {
  using C = rrdb::Cfg;
  RrdbPlan* P = ctx->rrdb;
  CUtensorMap amap;
  int rc = make_map(ctx, &amap, in, n, H, W, in_cs, rrdb::ROWS + 2, L.nc * 64 < in_cs && ctx->opt_rrdb_promo128);
  if (rc) return rc;
  const bool half_last = L.cin % 64 == 32 && ctx->opt_rrdb_half;
  CUtensorMap hmap = amap;
  if (half_last && (rc = make_map(ctx, &hmap, in, n, H, W, in_cs, rrdb::ROWS + 2, true, true))) return rc;
  const unsigned abit = 1u;
  if (!(P->attr_mask & abit)) {
    VEX_CUDA(ctx, cudaFuncSetAttribute(rrdb::rconv3x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_MAX));
    P->attr_mask |= abit;
  }
  rrdb::Params p;
  p.wpack = L.wpack; p.bias = L.bias;
  p.nblk = L.cout / rrdb::CO;
  p.out = e.out; p.out_cs = e.out_cs; p.out_c0 = e.out_c0;
  p.r1 = e.r1; p.r1_cs = e.r1_cs; p.a1 = e.a1;
  p.r2 = e.r2; p.r2_cs = e.r2_cs; p.a2 = e.a2;
  p.lrelu = e.lrelu; p.ups2 = e.ups2;
  p.H = H; p.W = W; p.NC = L.nc;
  p.sa = C::stages(L.nc * 3);                           // what the layer's weights leave of the 227 KB
  if (ctx->opt_rrdb_stages >= 2 && ctx->opt_rrdb_stages < p.sa) p.sa = ctx->opt_rrdb_stages;   // A/B against fewer
  p.klast = L.cin % 64 == 32 ? 2 : 4;
  p.half_last = half_last ? 1 : 0;
  p.tiles_x = (W + rrdb::TILE_W - 1) / rrdb::TILE_W;
  p.tiles_y = (H + rrdb::ROWS - 1) / rrdb::ROWS;
  p.n_img = n;
  p.dbg = nullptr;
  p.probe = ctx->opt_thin_probe;
  p.reverse = ctx->opt_rrdb_zigzag ? (int)(P->launch_parity++ & 1u) : 0;
  long long total = (long long)p.tiles_x * p.tiles_y * n;
  p.n_tab = 0;
  memset(p.vh, 0, sizeof(p.vh)); memset(p.vw, 0, sizeof(p.vw)); memset(p.item0, 0, sizeof(p.item0));
  if (!P->canvas_dims.empty()) {
    if ((int)P->canvas_dims.size() != 2 * n || n > rrdb::MAXB || P->canvas_h0 <= 0 || H % P->canvas_h0 || H > 32767 || W > 32767)
      return fail(ctx, VEX_ERR_INVALID, "rrdb conv: inconsistent canvas table");
    const int f = H / P->canvas_h0;
    p.n_tab = n;
    total = 0;
    for (int i = 0; i < n; ++i) {
      p.vh[i] = (short)(P->canvas_dims[2 * i] * f); p.vw[i] = (short)(P->canvas_dims[2 * i + 1] * f);
      p.item0[i] = (int)total;
      total += (long long)rrdb::items_x(p.vw[i], W) * rrdb::items_y(p.vh[i], H);
    }
    p.item0[n] = (int)total;
  }
  total *= p.nblk;                                      // work items = (tile, half)
  if (total > 0x7fffffffLL) return fail(ctx, VEX_ERR_INVALID, "rrdb conv: too many tiles");
  if (ctx->opt_conv_dbg) {
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&p.dbg), (size_t)ctx->sm_count * 64));
    VEX_CUDA(ctx, cudaMemsetAsync(p.dbg, 0, (size_t)ctx->sm_count * 64, s));
  }
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = ctx->opt_pdl ? 1 : 0;
  unsigned grid_x = (unsigned)std::min<long long>(total, ctx->sm_count);
  if (p.nblk == 2) grid_x &= ~1u;                       // even grid: a CTA's parity is the half it works on (total is even)
  cfg.gridDim = dim3(grid_x);
  cfg.blockDim = dim3(rrdb::NTHREADS);
  cfg.dynamicSmemBytes = (size_t)C::smem(p.sa, L.nc * 3);
  cfg.stream = s;
  cfg.attrs = attr; cfg.numAttrs = 1;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (p.dbg) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, s); }
  VEX_CUDA(ctx, cudaLaunchKernelEx(&cfg, rrdb::rconv3x3_kernel, amap, hmap, p));
  ctx->launches++;
  VEX_CUDA(ctx, cudaGetLastError());
  if (p.dbg) {
    cudaEventRecord(e1, s);
    const unsigned grid = cfg.gridDim.x;
    std::vector<long long> h((size_t)grid * 8);
    VEX_CUDA(ctx, cudaStreamSynchronize(s));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    VEX_CUDA(ctx, cudaMemcpy(h.data(), p.dbg, h.size() * 8, cudaMemcpyDeviceToHost));
    cudaFree(p.dbg);
    double a[8] = {0};
    for (unsigned b = 0; b < grid; ++b)
      for (int i = 0; i < 8; ++i) a[i] += (double)h[(size_t)b * 8 + i] / grid;
    const double items = (double)total / grid;
    const double mma_clk = ((double)(L.nc - 1) * 4 + p.klast) * 3 * rrdb::APS;   // MMAs (N = 96) per item
    fprintf(stderr, "[rconv_dbg cin=%d cout=%d nblk=%d %dx%d] %.1f us, items/cta=%.2f, cycles/item: MMA total=%.0f (%.1f clk/MMA) wait_acc_empty=%.0f "
                    "wait_a_full=%.0f wait_w_full=%.0f | epilogue wait_acc_full=%.0f work=%.0f | producer wait_a_empty=%.0f\n",
            L.cin, L.cout, p.nblk, H, W, ms * 1e3, items, a[0] / items, a[0] / items / mma_clk, a[1] / items, a[2] / items, a[3] / items,
            a[4] / items, a[5] / items, a[6] / items);
  }
  return VEX_OK;
}

int run_rconv(vex_ctx* ctx, const RConv& L, const __half* in, int in_cs, const Epi& e, int n, int H, int W,
              cudaStream_t s) {
  return launch_rconv(ctx, L, in, in_cs, e, n, H, W, s);
}

// fp16 NHWC channel slice -> fp32 NCHW (parity taps).  One thread per (pixel, channel).
__global__ void slice_f16_to_nchw_f32_kernel(const __half* __restrict__ in, float* __restrict__ out, size_t npix, int HW,
                                             int cs, int c0, int C) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix * C) return;
  const int c = (int)(i % C);
  const size_t pix = i / C;
  const size_t n = pix / HW, hw = pix % HW;
  out[(n * C + c) * (size_t)HW + hw] = __half2float(in[pix * cs + c0 + c]);
}

// canvas batching: conv_first runs over whole canvases, so what it wrote outside an image's valid part (its bias) is
// cleared before the trunk reads it as padding.  One thread per (pixel, 16-byte chunk of the 64 fp16 channels).
struct CanvasTab { short vh[rrdb::MAXB], vw[rrdb::MAXB]; };
__global__ void zero_outside_kernel(__half* __restrict__ buf, size_t nchunk, int H, int W, CanvasTab t) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nchunk) return;
  const size_t pix = i >> 3;
  const int x = (int)(pix % W), y = (int)((pix / W) % H), n = (int)(pix / ((size_t)W * H));
  if (y >= t.vh[n] || x >= t.vw[n]) reinterpret_cast<uint4*>(buf)[i] = make_uint4(0, 0, 0, 0);
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Layout {
  size_t d[3], feat, u1, u2, u3, total;
};
Layout ws_layout(const vex_rrdb_desc& a, size_t npix) {
  Layout L;
  size_t off = 0;
  for (int i = 0; i < 3; ++i) { L.d[i] = off; off += align_up(npix * 192 * 2, 1024); }
  L.feat = off; off += align_up(npix * 64 * 2, 1024);
  const size_t s2 = 4, sq = (size_t)a.scale * a.scale;
  L.u1 = off; off += align_up(npix * s2 * 64 * 2, 1024);          // nearest-2x of the trunk: conv_up1's input
  L.u2 = off; off += align_up(npix * sq * 64 * 2, 1024);          // conv_up1's output (scale 4: replicated again); later conv_hr's output
  L.u3 = off; off += align_up(npix * sq * 64 * 2, 1024);          // conv_up2's output (scale 4) / unused
  L.total = off;
  return L;
}

}  // namespace

// ------------------------------------------------------------------------------------------- weights

int rrdb_load_weights(vex_ctx* ctx, const vex_rrdb_desc* arch, const float* const* t, int n_tensors) {
  if (!arch || !t) return fail(ctx, VEX_ERR_INVALID, "vex_rrdb_load_weights: null argument");
  if (arch->num_feat != 64 || arch->num_grow_ch != 32 || (arch->scale != 2 && arch->scale != 4) || arch->num_block < 1 ||
      arch->num_block > 64)
    return fail(ctx, VEX_ERR_UNSUPPORTED,
                "vex_rrdb_load_weights: num_feat 64, num_grow_ch 32 (the handler's fixed values, realesrgan_handler.py:148-150), scale 2 or 4");
  const int n_up = arch->scale == 4 ? 2 : 1;
  const int want = 2 + arch->num_block * 30 + 2 + 2 * n_up + 4;
  if (n_tensors != want)
    return fail(ctx, VEX_ERR_INVALID, "vex_rrdb_load_weights: expected " + std::to_string(want) + " tensors in state_dict order");
  for (int i = 0; i < n_tensors; ++i)
    if (!t[i]) return fail(ctx, VEX_ERR_INVALID, "vex_rrdb_load_weights: null tensor");
  if (!ctx->rrdb) ctx->rrdb = new RrdbPlan();
  RrdbPlan* P = ctx->rrdb;
  P->arch = *arch;
  if (!P->encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    VEX_CUDA(ctx, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn || q != cudaDriverEntryPointSuccess)
      return fail(ctx, VEX_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled not available in this driver");
    P->encode = fn;
  }
  int rc, i = 0;
  {  // conv_first [64][3][3][3]: im2col27 image, rows = output channels; k = (ky*3+kx)*3+ci holds bf16(w/255), k + 32 the
     // part that rounding dropped; k = 27, 28 the bias as hi + lo bf16 (the builders put 1.0 there)
    const float *w = t[i], *b = t[i + 1];
    i += 2;
    std::vector<uint16_t> img((size_t)64 * 64, 0);
    for (int r = 0; r < 64; ++r)
      for (int k = 0; k < 64; ++k) {
        float v = 0.f;
        const int kk = k >= 32 ? k - 32 : k;
        if (k == 27 || k == 28) {
          const float hi = bf16_to_f32(f32_to_bf16(b[r]));
          v = k == 27 ? hi : b[r] - hi;
        } else if (kk < 27 && k != 59 && k != 60) {
          const int ci = kk % 3, tap = kk / 3;
          const float full = w[((size_t)r * 3 + ci) * 9 + tap] * (1.0f / 255.0f);   // the /255.0 of :311
          const float hi = bf16_to_f32(f32_to_bf16(full));
          v = k < 32 ? hi : full - hi;
        }
        img[(size_t)r * 64 + (((k >> 3) ^ (r & 7)) << 3) + (k & 7)] = f32_to_bf16(v);
      }
    if ((rc = upload(ctx, &P->first_w, img.data(), img.size() * 2))) return rc;
  }
  P->rdb.resize((size_t)arch->num_block * 15);
  for (int b = 0; b < arch->num_block; ++b)
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 5; ++c) {
        RConv& L = P->rdb[((size_t)b * 3 + r) * 5 + c];
        L.cin = 64 + 32 * c;
        L.cout = c == 4 ? 64 : 32;
        if ((rc = pack_rconv(ctx, L, t[i], t[i + 1]))) return rc;
        i += 2;
      }
  RConv* tail[4] = {&P->body, &P->up1, n_up == 2 ? &P->up2 : nullptr, &P->hr};
  for (RConv* L : tail) {
    if (!L) continue;
    L->cin = 64; L->cout = 64;
    if ((rc = pack_rconv(ctx, *L, t[i], t[i + 1]))) return rc;
    i += 2;
  }
  {  // conv_last [3][64][3][3] as a 1x1 GEMM to 27 (tap, co) columns: row n = tap*3+co (27 of 32), K = input channel, fp16
    const float *w = t[i], *b = t[i + 1];
    std::vector<uint16_t> img((size_t)32 * 64, 0);
    for (int r = 0; r < 27; ++r)
      for (int k = 0; k < 64; ++k)
        img[(size_t)r * 64 + (((k >> 3) ^ (r & 7)) << 3) + (k & 7)] = f32_to_f16(w[((size_t)(r % 3) * 64 + k) * 9 + r / 3]);
    if ((rc = upload(ctx, &P->last_w, img.data(), img.size() * 2))) return rc;
    if ((rc = upload(ctx, &P->last_b, b, 12))) return rc;
  }
  return VEX_OK;
}

// ------------------------------------------------------------------------------------------- forward

size_t rrdb_workspace_bytes(const vex_ctx* ctx, int n, int h, int w) {
  if (!ctx->rrdb) return 0;
  return ws_layout(ctx->rrdb->arch, (size_t)n * h * w).total;
}

// tap: 0 = the network (out_u8 and/or out_f32 = clamp(conv_last) NCHW [n][3][sH][sW]); 1 = conv_first; 2 = feat + body_feat;
// 3 = lrelu(conv_hr); 10 + i = output of body block i; 1000 + c (c = 1..4) = x_c of body[0].rdb1, 1005 = body[0].rdb1's output
int rrdb_forward(vex_ctx* ctx, const uint8_t* in, uint8_t* out_u8, float* out_f32, int tap, int n, int H, int W,
                 int in_bgr, int out_bgr, cudaStream_t s) {
  RrdbPlan* P = ctx->rrdb;
  if (!P || !P->first_w) return fail(ctx, VEX_ERR_NO_WEIGHTS, "vex_rrdb_load_weights has not been called");
  if ((long long)H * W * P->arch.scale * P->arch.scale > (1LL << 30)) return fail(ctx, VEX_ERR_INVALID, "frame too large");
  const vex_rrdb_desc& a = P->arch;
  const size_t npix = (size_t)n * H * W;
  const Layout L = ws_layout(a, npix);
  P->canvas_h0 = H;
  if (L.total > P->ws_bytes) {
    VEX_CUDA(ctx, cudaDeviceSynchronize());
    if (P->ws) cudaFree(P->ws);
    P->ws = nullptr; P->ws_bytes = 0;
    VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&P->ws), L.total));
    P->ws_bytes = L.total;
    // every value a conv reads must be finite, including the not-yet-written slices it multiplies by zero weights
    VEX_CUDA(ctx, cudaMemsetAsync(P->ws, 0, L.total, s));
  }
  // one workspace per plan: order forwards issued on different streams (same rule as the CNN's, vex_internal.cuh)
  if (!ctx->ev_ws_rrdb) VEX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_ws_rrdb, cudaEventDisableTiming));
  if (ctx->ws_rrdb_used) VEX_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_ws_rrdb, 0));
  struct WsMark {
    vex_ctx* c; cudaStream_t st;
    ~WsMark() { if (cudaEventRecord(c->ev_ws_rrdb, st) == cudaSuccess) c->ws_rrdb_used = true; }
  } ws_mark{ctx, s};

  __half* D[3] = {reinterpret_cast<__half*>(P->ws + L.d[0]), reinterpret_cast<__half*>(P->ws + L.d[1]),
                  reinterpret_cast<__half*>(P->ws + L.d[2])};
  __half* feat = reinterpret_cast<__half*>(P->ws + L.feat);
  __half* u1 = reinterpret_cast<__half*>(P->ws + L.u1);
  __half* u2 = reinterpret_cast<__half*>(P->ws + L.u2);
  __half* u3 = reinterpret_cast<__half*>(P->ws + L.u3);
  const int HW = H * W;
  auto emit = [&](const __half* buf, int cs, int c0, int C, size_t px, int hw) -> int {
    slice_f16_to_nchw_f32_kernel<<<(unsigned)((px * C + 255) / 256), 256, 0, s>>>(buf, out_f32, px, hw, cs, c0, C);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    return VEX_OK;
  };
  int rc;
  {  // conv_first -> feat, and a copy of it into the first dense buffer's x slice
    using C = thin::I27Cfg<64>;
    if (!(P->attr_mask & 4u)) {
      VEX_CUDA(ctx, cudaFuncSetAttribute(thin::im2col27_kernel<thin::MODE_FIRST64, 64>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
      P->attr_mask |= 4u;
    }
    thin::I27Params p;
    p.in = in; p.wpack = P->first_w; p.out = reinterpret_cast<__nv_bfloat16*>(feat); p.H = H; p.W = W;
    p.npix = (long long)npix; p.in_bgr = in_bgr; p.probe = 0; p.out_cs = 64; p.dbg = nullptr;
    const long long items = (p.npix + 127) / 128;
    thin::im2col27_kernel<thin::MODE_FIRST64, 64><<<(unsigned)std::min<long long>(items, ctx->sm_count), thin::I27_THREADS,
                                                    C::SMEM, s>>>(p);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
    if (!P->canvas_dims.empty()) {
      if ((int)P->canvas_dims.size() != 2 * n || n > rrdb::MAXB) return fail(ctx, VEX_ERR_INVALID, "rrdb: inconsistent canvas table");
      CanvasTab t{};
      for (int i = 0; i < n; ++i) { t.vh[i] = (short)P->canvas_dims[2 * i]; t.vw[i] = (short)P->canvas_dims[2 * i + 1]; }
      const size_t nchunk = npix * 8;
      zero_outside_kernel<<<(unsigned)((nchunk + 255) / 256), 256, 0, s>>>(feat, nchunk, H, W, t);
      ctx->launches++;
      VEX_CUDA(ctx, cudaGetLastError());
    }
    VEX_CUDA(ctx, cudaMemcpy2DAsync(D[0], 384, feat, 128, 128, npix, cudaMemcpyDeviceToDevice, s));
  }
  if (out_f32 && tap == 1) return emit(feat, 64, 0, 64, npix, HW);
  int ia = 0, ib = 1, ic = 2;
  for (int b = 0; b < a.num_block; ++b) {
    const int order[3][3] = {{ia, ib, -1}, {ib, ic, -1}, {ic, ib, ia}};   // {input buffer, output buffer, RRDB residual or -1}
    for (int r = 0; r < 3; ++r) {
      __half* X = D[order[r][0]];
      const RConv* Ls = &P->rdb[((size_t)b * 3 + r) * 5];
      for (int c = 0; c < 4; ++c) {                                        // x_{c+1} = lrelu(conv(cat(x, x1..xc)))  (:50-53)
        Epi e; e.out = X; e.out_cs = 192; e.out_c0 = 64 + 32 * c; e.lrelu = 1;
        if ((rc = run_rconv(ctx, Ls[c], X, 192, e, n, H, W, s))) return rc;
        if (out_f32 && b == 0 && r == 0 && tap == 1001 + c) return emit(X, 192, 64 + 32 * c, 32, npix, HW);
      }
      Epi e; e.out = D[order[r][1]]; e.out_cs = 192; e.out_c0 = 0;
      e.r1 = X; e.r1_cs = 192; e.a1 = 0.2f;                                 // x5 * 0.2 + x  (:57)
      if (order[r][2] >= 0) { e.r2 = D[order[r][2]]; e.r2_cs = 192; e.a2 = 0.2f; }   // out * 0.2 + x  (:36)
      if ((rc = run_rconv(ctx, Ls[4], X, 192, e, n, H, W, s))) return rc;
      if (out_f32 && b == 0 && r == 0 && tap == 1005) return emit(e.out, 192, 0, 64, npix, HW);
    }
    const int na = ib, nb = ic, nc = ia;
    ia = na; ib = nb; ic = nc;
    if (out_f32 && tap == 10 + b) return emit(D[ia], 192, 0, 64, npix, HW);
  }
  {  // conv_body, + feat, replicated 2x2 into u1 (:88-89, :93)
    Epi e; e.out = u1; e.out_cs = 64; e.out_c0 = 0; e.r1 = feat; e.r1_cs = 64; e.a1 = 1.0f; e.ups2 = 1;
    if (out_f32 && tap == 2) { e.out = D[ib]; e.out_cs = 192; e.ups2 = 0; }
    if ((rc = run_rconv(ctx, P->body, D[ia], 192, e, n, H, W, s))) return rc;
    if (out_f32 && tap == 2) return emit(D[ib], 192, 0, 64, npix, HW);
  }
  const __half* hr_in;
  int Hh, Wh;
  {  // conv_up1 (+ conv_up2 for scale 4) with LeakyReLU (:93-96)
    Epi e; e.out = u2; e.out_cs = 64; e.out_c0 = 0; e.lrelu = 1; e.ups2 = a.scale == 4;
    if ((rc = run_rconv(ctx, P->up1, u1, 64, e, n, 2 * H, 2 * W, s))) return rc;
    if (a.scale == 4) {
      Epi e2; e2.out = u3; e2.out_cs = 64; e2.out_c0 = 0; e2.lrelu = 1;
      if ((rc = run_rconv(ctx, P->up2, u2, 64, e2, n, 4 * H, 4 * W, s))) return rc;
      hr_in = u3; Hh = 4 * H; Wh = 4 * W;
    } else {
      hr_in = u2; Hh = 2 * H; Wh = 2 * W;
    }
  }
  __half* hr_out = a.scale == 4 ? u2 : u3;
  {
    Epi e; e.out = hr_out; e.out_cs = 64; e.out_c0 = 0; e.lrelu = 1;      // lrelu(conv_hr(feat))  (:98)
    if ((rc = run_rconv(ctx, P->hr, hr_in, 64, e, n, Hh, Wh, s))) return rc;
    if (out_f32 && tap == 3) return emit(hr_out, 64, 0, 64, (size_t)n * Hh * Wh, Hh * Wh);
  }
  {  // conv_last + clamp + x255 + truncating u8 (:99, :318-319)
    if (!(P->attr_mask & 8u)) {
      VEX_CUDA(ctx, cudaFuncSetAttribute(thin::gemm27_kernel<thin::MODE_LAST_F16>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, thin::G27Cfg::SMEM));
      P->attr_mask |= 8u;
    }
    CUtensorMap amap;
    if ((rc = make_map(ctx, &amap, hr_out, n, Hh, Wh, 64, 8))) return rc;
    thin::G27Params p;
    p.wpack = P->last_w; p.bias = P->last_b; p.out_f4 = nullptr; p.out_u8 = out_u8; p.out_nchw = out_f32;
    p.H = Hh; p.W = Wh; p.NC = 1;
    p.tiles_x = (Wh + thin::G27_TW - 1) / thin::G27_TW;
    p.tiles_y = (Hh + thin::G27_ROWS - 1) / thin::G27_ROWS;
    p.n_img = n; p.out_bgr = out_bgr;
    const long long total = (long long)p.tiles_x * p.tiles_y * n;
    if (total > 0x7fffffffLL) return fail(ctx, VEX_ERR_INVALID, "rrdb conv_last: too many tiles");
    thin::gemm27_kernel<thin::MODE_LAST_F16><<<(unsigned)std::min<long long>(total, ctx->sm_count), thin::G27_THREADS,
                                               thin::G27Cfg::SMEM, s>>>(amap, p);
    ctx->launches++;
    VEX_CUDA(ctx, cudaGetLastError());
  }
  return VEX_OK;
}

// `_enhance_frame` (:286-306) on the device: one frame, tiled like `_tile_process` (:324-372) when either side exceeds
// tile_size, else `_process_single`.  in / out are packed u8 HWC device buffers.
//
// The reference runs its tiles one after the other.  Here the tiles of a frame are independent images of one BATCH: each
// is cropped (with its padding) to the top-left corner of its own Hc x Wc canvas, where Hc x Wc is the largest padded tile,
// and every layer is launched once for all of them (option rrdb_batch, default on).  The conv kernel knows each canvas's
// valid size and keeps everything outside it zero, so a tile sees the same zero padding as when it runs alone and the
// result is bit-identical (tests/test_rrdb_gpu.py) - but a layer launch now has ~12x the work items: the fill/drain of the
// per-CTA pipelines, the launch gap and the rounding-up of items to whole waves of 148 CTAs are paid once per layer
// instead of once per layer AND tile.  Batches hold at most MAXB tiles and ws_budget_mb of activations.
int rrdb_frame(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int h, int w, int tile_size, int tile_pad, int in_bgr,
               int out_bgr, cudaStream_t s) {
  RrdbPlan* P = ctx->rrdb;
  if (!P || !P->first_w) return fail(ctx, VEX_ERR_NO_WEIGHTS, "vex_rrdb_load_weights has not been called");
  if (tile_size <= 0 || tile_pad < 0) return fail(ctx, VEX_ERR_INVALID, "vex_rrdb_frame_u8: bad tile geometry");
  const int S = P->arch.scale;
  if (h <= tile_size && w <= tile_size) return rrdb_forward(ctx, in, out, nullptr, 0, 1, h, w, in_bgr, out_bgr, s);

  struct Tile { int sx, sy, ex, ey, px0, py0, tw, th; };
  std::vector<Tile> tiles;
  const int tiles_x = (w + tile_size - 1) / tile_size, tiles_y = (h + tile_size - 1) / tile_size;
  for (int ty = 0; ty < tiles_y; ++ty)
    for (int tx = 0; tx < tiles_x; ++tx) {
      Tile t;
      t.sx = tx * tile_size; t.sy = ty * tile_size;
      t.ex = std::min(t.sx + tile_size, w); t.ey = std::min(t.sy + tile_size, h);
      t.px0 = std::max(t.sx - tile_pad, 0); t.py0 = std::max(t.sy - tile_pad, 0);
      t.tw = std::min(t.ex + tile_pad, w) - t.px0; t.th = std::min(t.ey + tile_pad, h) - t.py0;
      tiles.push_back(t);
    }
  const size_t per_px = ws_layout(P->arch, 1 << 20).total >> 20;            // workspace bytes per input pixel
  const size_t budget_px = std::max<size_t>(((size_t)ctx->opt_ws_budget_mb << 20) / std::max<size_t>(per_px, 1), 1);
  const int max_batch = std::min(std::max(ctx->opt_rrdb_batch, 1), rrdb::MAXB);

  size_t first = 0;
  while (first < tiles.size()) {
    // greedy batch of consecutive tiles: at most MAXB canvases, n * Hc * Wc within the workspace budget
    int Hc = tiles[first].th, Wc = tiles[first].tw;
    size_t last = first + 1;
    while (last < tiles.size() && (int)(last - first) < max_batch) {
      const int nh = std::max(Hc, tiles[last].th), nw = std::max(Wc, tiles[last].tw);
      if ((last - first + 1) * (size_t)nh * nw > budget_px) break;
      Hc = nh; Wc = nw; ++last;
    }
    const int n = (int)(last - first);
    const size_t canvas_in = (size_t)Hc * Wc * 3, canvas_out = canvas_in * S * S;
    const size_t need_in = canvas_in * n, need_out = canvas_out * n;
    if (need_in > P->tile_in_bytes || need_out > P->tile_out_bytes) {
      VEX_CUDA(ctx, cudaDeviceSynchronize());
      if (P->tile_in) cudaFree(P->tile_in);
      if (P->tile_out) cudaFree(P->tile_out);
      P->tile_in = P->tile_out = nullptr;
      P->tile_in_bytes = P->tile_out_bytes = 0;
      VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&P->tile_in), need_in));
      VEX_CUDA(ctx, cudaMalloc(reinterpret_cast<void**>(&P->tile_out), need_out));
      P->tile_in_bytes = need_in; P->tile_out_bytes = need_out;
    }
    bool uniform = true;
    for (size_t i = first; i < last; ++i) uniform = uniform && tiles[i].th == Hc && tiles[i].tw == Wc;
    if (!uniform) VEX_CUDA(ctx, cudaMemsetAsync(P->tile_in, 0, need_in, s));   // zero = what a smaller tile has beyond its edge
    std::vector<int> dims;
    for (size_t i = first; i < last; ++i) {
      const Tile& t = tiles[i];
      VEX_CUDA(ctx, cudaMemcpy2DAsync(P->tile_in + (i - first) * canvas_in, (size_t)Wc * 3,
                                      in + ((size_t)t.py0 * w + t.px0) * 3, (size_t)w * 3, (size_t)t.tw * 3, t.th,
                                      cudaMemcpyDeviceToDevice, s));
      dims.push_back(t.th); dims.push_back(t.tw);
    }
    if (!uniform) P->canvas_dims = dims;
    const int rc = rrdb_forward(ctx, P->tile_in, P->tile_out, nullptr, 0, n, Hc, Wc, in_bgr, out_bgr, s);
    P->canvas_dims.clear();
    if (rc) return rc;
    for (size_t i = first; i < last; ++i) {
      const Tile& t = tiles[i];
      const size_t pl = (size_t)(t.sx - t.px0) * S, pt = (size_t)(t.sy - t.py0) * S;
      VEX_CUDA(ctx, cudaMemcpy2DAsync(out + ((size_t)t.sy * S * w * S + (size_t)t.sx * S) * 3, (size_t)w * S * 3,
                                      P->tile_out + (i - first) * canvas_out + (pt * Wc * S + pl) * 3, (size_t)Wc * S * 3,
                                      (size_t)(t.ex - t.sx) * S * 3, (size_t)(t.ey - t.sy) * S, cudaMemcpyDeviceToDevice, s));
    }
    first = last;
  }
  return VEX_OK;
}

void rrdb_destroy(vex_ctx* ctx) {
  RrdbPlan* P = ctx->rrdb;
  if (!P) return;
  auto freec = [](RConv& L) { if (L.wpack) cudaFree(L.wpack); if (L.bias) cudaFree(L.bias); };
  for (RConv& L : P->rdb) freec(L);
  freec(P->body); freec(P->up1); freec(P->up2); freec(P->hr);
  if (P->first_w) cudaFree(P->first_w);
  if (P->last_w) cudaFree(P->last_w);
  if (P->last_b) cudaFree(P->last_b);
  if (P->ws) cudaFree(P->ws);
  if (P->tile_in) cudaFree(P->tile_in);
  if (P->tile_out) cudaFree(P->tile_out);
  delete P;
  ctx->rrdb = nullptr;
}

}  // namespace vex
