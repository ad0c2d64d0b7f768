// C ABI of libvex.so (include/vex.h): context, argument checking, host-buffer streaming.
#include <algorithm>
#include <string.h>
#include <vector>

#include "vex_internal.cuh"

namespace {
std::string g_create_error;

int ensure_slots(vex_ctx* ctx, size_t in_bytes, size_t out_bytes) {
  if (!ctx->st_h2d) {
    VEX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->st_h2d, cudaStreamNonBlocking));
    VEX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->st_comp, cudaStreamNonBlocking));
    VEX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->st_d2h, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      VEX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_h2d[i], cudaEventDisableTiming));
      VEX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_comp[i], cudaEventDisableTiming));
      VEX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_d2h[i], cudaEventDisableTiming));
    }
  }
  if (in_bytes > ctx->slot_in_bytes || out_bytes > ctx->slot_out_bytes) {
    VEX_CUDA(ctx, cudaDeviceSynchronize());
    for (int i = 0; i < 2; ++i) {
      if (ctx->slot_in[i]) cudaFree(ctx->slot_in[i]);
      if (ctx->slot_out[i]) cudaFree(ctx->slot_out[i]);
      ctx->slot_in[i] = ctx->slot_out[i] = nullptr;
    }
    ctx->slot_in_bytes = ctx->slot_out_bytes = 0;
    for (int i = 0; i < 2; ++i) {
      VEX_CUDA(ctx, cudaMalloc(&ctx->slot_in[i], in_bytes));
      VEX_CUDA(ctx, cudaMalloc(&ctx->slot_out[i], out_bytes));
    }
    ctx->slot_in_bytes = in_bytes;
    ctx->slot_out_bytes = out_bytes;
  }
  return VEX_OK;
}
}  // namespace

extern "C" {

int vex_abi_version(void) { return VEX_ABI_VERSION; }

int vex_create(int device_ordinal, vex_ctx** out) {
  if (!out) return VEX_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    g_create_error = std::string("vex_create: no CUDA device (") +
                     (e != cudaSuccess ? cudaGetErrorString(e) : "count=0") +
                     "); libvex has no CPU fallback";
    return VEX_ERR_CUDA;
  }
  if (device_ordinal < 0 || device_ordinal >= count) {
    g_create_error = "vex_create: device ordinal out of range";
    return VEX_ERR_INVALID;
  }
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device_ordinal)) != cudaSuccess) {
    g_create_error = std::string("vex_create: ") + cudaGetErrorString(e);
    return VEX_ERR_CUDA;
  }
  if (prop.major != 10) {
    g_create_error = "vex_create: device is sm_" + std::to_string(prop.major * 10 + prop.minor) +
                     "; libvex is built for sm_100a only";
    return VEX_ERR_UNSUPPORTED;
  }
  if ((e = cudaSetDevice(device_ordinal)) != cudaSuccess) {
    g_create_error = std::string("vex_create: cudaSetDevice: ") + cudaGetErrorString(e);
    return VEX_ERR_CUDA;
  }
  vex_ctx* ctx = new vex_ctx();
  ctx->device = device_ordinal;
  ctx->sm_count = prop.multiProcessorCount;
  *out = ctx;
  return VEX_OK;
}

int vex_destroy(vex_ctx* ctx) {
  if (!ctx) return VEX_ERR_INVALID;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  vex::cnn_destroy(ctx);
  if (ctx->bilateral_cw) cudaFree(ctx->bilateral_cw);
  for (int i = 0; i < 2; ++i) {
    if (ctx->slot_in[i]) cudaFree(ctx->slot_in[i]);
    if (ctx->slot_out[i]) cudaFree(ctx->slot_out[i]);
    if (ctx->ev_h2d[i]) cudaEventDestroy(ctx->ev_h2d[i]);
    if (ctx->ev_comp[i]) cudaEventDestroy(ctx->ev_comp[i]);
    if (ctx->ev_d2h[i]) cudaEventDestroy(ctx->ev_d2h[i]);
  }
  if (ctx->st_h2d) cudaStreamDestroy(ctx->st_h2d);
  if (ctx->st_comp) cudaStreamDestroy(ctx->st_comp);
  if (ctx->st_d2h) cudaStreamDestroy(ctx->st_d2h);
  delete ctx;
  return VEX_OK;
}

const char* vex_last_error(const vex_ctx* ctx) {
  return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

int vex_bicubic2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w,
                     int round_mode, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::launch_bicubic2x(ctx, in_hwc, out_hwc, n, h, w, round_mode, (cudaStream_t)stream);
}

int vex_lanczos4_2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::launch_lanczos4_2x(ctx, in_hwc, out_hwc, n, h, w, (cudaStream_t)stream);
}

int vex_resize2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, int interpolation,
                    void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  switch (interpolation) {                     // OpenCV's InterpolationFlags values
    case 0: case 3: return vex::launch_resize2x(ctx, in_hwc, out_hwc, n, h, w, 0, (cudaStream_t)stream);
    case 1: return vex::launch_resize2x(ctx, in_hwc, out_hwc, n, h, w, 1, (cudaStream_t)stream);
    case 2: return vex::launch_bicubic2x(ctx, in_hwc, out_hwc, n, h, w, 1, (cudaStream_t)stream);
    case 4: return vex::launch_lanczos4_2x(ctx, in_hwc, out_hwc, n, h, w, (cudaStream_t)stream);
    default: break;
  }
  return vex::fail(ctx, VEX_ERR_INVALID,
                   "vex_resize2x_u8: interpolation must be 0 (NEAREST), 1 (LINEAR), 2 (CUBIC), 3 (AREA) or 4 (LANCZOS4)");
}

int vex_bilateral5_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, float sigma_color,
                      float sigma_space, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::launch_bilateral5(ctx, in_hwc, out_hwc, n, h, w, sigma_color, sigma_space, (cudaStream_t)stream);
}

int vex_filter3x3_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, const float* kernel9,
                     void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::launch_filter3x3(ctx, in_hwc, out_hwc, n, h, w, kernel9, (cudaStream_t)stream);
}

int vex_cnn_load_weights(vex_ctx* ctx, const vex_cnn_desc* arch, const float* const* tensors,
                         int n_tensors) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::cnn_load_weights(ctx, arch, tensors, n_tensors);
}

int vex_cnn2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w,
                 int in_bgr, int out_bgr, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_hwc || n <= 0 || h <= 0 || w <= 0)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_cnn2x_u8: bad argument");
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::cnn_forward(ctx, in_hwc, out_hwc, nullptr, 0, n, h, w, in_bgr != 0, out_bgr != 0,
                          (cudaStream_t)stream);
}

int vex_cnn2x_f32_debug(vex_ctx* ctx, const uint8_t* in_hwc, float* out_nchw, int n, int h, int w,
                        int in_bgr, int tap, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_nchw || n <= 0 || h <= 0 || w <= 0 || tap < 0 || tap > 6 + VEX_CNN_MAX_EXTRA)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_cnn2x_f32_debug: bad argument");
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::cnn_forward(ctx, in_hwc, nullptr, out_nchw, tap, n, h, w, in_bgr != 0, 0,
                          (cudaStream_t)stream);
}

int vex_workspace_bytes(const vex_ctx* ctx, int n, int h, int w, size_t* out) {
  if (!ctx || !out || n <= 0 || h <= 0 || w <= 0) return VEX_ERR_INVALID;
  *out = vex::cnn_workspace_bytes(ctx, n, h, w);
  return VEX_OK;
}

int vex_host_alloc(vex_ctx* ctx, size_t bytes, void** out) {
  if (!ctx || !out || bytes == 0) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  VEX_CUDA(ctx, cudaHostAlloc(out, bytes, cudaHostAllocDefault));
  return VEX_OK;
}

int vex_host_free(vex_ctx* ctx, void* p) {
  if (!ctx || !p) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaFreeHost(p));
  return VEX_OK;
}

int vex_upscale_host(vex_ctx* ctx, int mode, const uint8_t* in_host, uint8_t* out_host, int n,
                     int h, int w, int chunk, int flags) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_host || !out_host || n <= 0 || h <= 0 || w <= 0 || chunk <= 0 ||
      (mode != VEX_MODE_BICUBIC && mode != VEX_MODE_CNN))
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_upscale_host: bad argument");
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  chunk = std::min(chunk, n);
  const size_t fin = (size_t)h * w * 3, fout = fin * 4;
  int rc = ensure_slots(ctx, fin * chunk, fout * chunk);
  if (rc) return rc;
  // Three-stage pipeline over two device slots: chunk k's H2D (copy-in stream), kernels (compute stream)
  // and D2H (copy-out stream) are chained by events, so H2D(k+1) and D2H(k-1) overlap compute(k), and the
  // CNN workspace is only ever touched from the compute stream.
  // Chunk schedule: the first H2D and the last D2H are the only copies no kernel hides, so long calls start and end with
  // a quarter-size chunk (48 frames at chunk 8: 2, 8, 8, 8, 8, 8, 4, 2).
  std::vector<int> sizes;
  {
    const int edge = (n >= 3 * chunk && chunk >= 4) ? chunk / 4 : 0;
    int mid = n - 2 * edge;
    if (edge) sizes.push_back(edge);
    while (mid > 0) { sizes.push_back(std::min(chunk, mid)); mid -= sizes.back(); }
    if (edge) sizes.push_back(edge);
  }
  int k = 0, f0 = 0;
  for (size_t ci = 0; ci < sizes.size(); f0 += sizes[ci], ++ci, ++k) {
    const int nf = sizes[ci], sl = k & 1;
    if (k >= 2) VEX_CUDA(ctx, cudaStreamWaitEvent(ctx->st_h2d, ctx->ev_comp[sl], 0));   // input slot consumed
    VEX_CUDA(ctx, cudaMemcpyAsync(ctx->slot_in[sl], in_host + fin * f0, fin * nf, cudaMemcpyHostToDevice,
                                  ctx->st_h2d));
    VEX_CUDA(ctx, cudaEventRecord(ctx->ev_h2d[sl], ctx->st_h2d));
    VEX_CUDA(ctx, cudaStreamWaitEvent(ctx->st_comp, ctx->ev_h2d[sl], 0));
    if (k >= 2) VEX_CUDA(ctx, cudaStreamWaitEvent(ctx->st_comp, ctx->ev_d2h[sl], 0));  // output slot drained
    if (mode == VEX_MODE_BICUBIC)
      rc = vex::launch_bicubic2x(ctx, ctx->slot_in[sl], ctx->slot_out[sl], nf, h, w, flags & 1, ctx->st_comp);
    else
      rc = vex::cnn_forward(ctx, ctx->slot_in[sl], ctx->slot_out[sl], nullptr, 0, nf, h, w, flags & 1,
                            (flags >> 1) & 1, ctx->st_comp);
    if (rc) {
      cudaDeviceSynchronize();
      return rc;
    }
    VEX_CUDA(ctx, cudaEventRecord(ctx->ev_comp[sl], ctx->st_comp));
    VEX_CUDA(ctx, cudaStreamWaitEvent(ctx->st_d2h, ctx->ev_comp[sl], 0));
    VEX_CUDA(ctx, cudaMemcpyAsync(out_host + fout * f0, ctx->slot_out[sl], fout * nf, cudaMemcpyDeviceToHost,
                                  ctx->st_d2h));
    VEX_CUDA(ctx, cudaEventRecord(ctx->ev_d2h[sl], ctx->st_d2h));
  }
  VEX_CUDA(ctx, cudaStreamSynchronize(ctx->st_d2h));
  VEX_CUDA(ctx, cudaStreamSynchronize(ctx->st_comp));
  VEX_CUDA(ctx, cudaStreamSynchronize(ctx->st_h2d));
  return VEX_OK;
}

int vex_cnn_profile_ms(vex_ctx* ctx, float* ms7) {
  if (!ctx || !ms7) return VEX_ERR_INVALID;
  VEX_CUDA(ctx, cudaSetDevice(ctx->device));
  return vex::cnn_profile_ms(ctx, ms7);
}

int vex_kernel_launches(const vex_ctx* ctx, unsigned long long* out) {
  if (!ctx || !out) return VEX_ERR_INVALID;
  *out = ctx->launches;
  return VEX_OK;
}

int vex_set_option(vex_ctx* ctx, const char* name, int value) {
  if (!ctx || !name) return VEX_ERR_INVALID;
  int* slot = nullptr;
  if (!strcmp(name, "desc_base_offset")) slot = &ctx->opt_desc_base_offset;
  else if (!strcmp(name, "force_generic_bicubic")) slot = &ctx->opt_force_generic_bicubic;
  else if (!strcmp(name, "force_generic_resize")) slot = &ctx->opt_force_generic_resize;
  else if (!strcmp(name, "bicubic_rows")) slot = &ctx->opt_bicubic_rows;
  else if (!strcmp(name, "conv_variant")) slot = &ctx->opt_conv_variant;
  else if (!strcmp(name, "thin_probe")) slot = &ctx->opt_thin_probe;
  else if (!strcmp(name, "conv3_bn")) slot = &ctx->opt_conv3_bn;
  else if (!strcmp(name, "ws_budget_mb")) slot = &ctx->opt_ws_budget_mb;
  else if (!strcmp(name, "profile")) slot = &ctx->opt_profile;
  else if (!strcmp(name, "thin_tc")) slot = &ctx->opt_thin_tc;
  else if (!strcmp(name, "fuse_tail")) slot = &ctx->opt_fuse_tail;
  else if (!strcmp(name, "conv_dbg")) slot = &ctx->opt_conv_dbg;
  else if (!strcmp(name, "epi_variant")) slot = &ctx->opt_epi_variant;
  if (!slot) return vex::fail(ctx, VEX_ERR_INVALID, std::string("vex_set_option: unknown ") + name);
  const int prev = *slot;
  *slot = value;
  if (slot == &ctx->opt_conv3_bn) {
    cudaSetDevice(ctx->device);
    if (int rc = vex::cnn_repack(ctx)) return rc;
  }
  return prev;
}

}  // extern "C"
