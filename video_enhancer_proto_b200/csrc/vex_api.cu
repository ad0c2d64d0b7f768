// C ABI of libvex.so (include/vex.h): context, argument checking, host-buffer streaming.
#include <algorithm>
#include <string.h>
#include <thread>
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

// Chunk sizes of a streaming call.  Only the first H2D and the last D2H of a call are copies no kernel hides (and, for
// vex_upscale_frames, the host staging of the first chunk), so a call starts with one frame, doubles up to `chunk`
// (the staging / H2D of chunk k+1 always fits under the kernels of chunk k) and ends with one frame:
// n = 8, chunk 8 -> 1, 2, 4, 1;  n = 48 -> 1, 2, 4, 8, 8, 8, 8, 8, 1.
std::vector<int> chunk_schedule(int n, int chunk) {
  std::vector<int> sizes;
  chunk = std::max(1, std::min(chunk, n));
  const int tail = n >= 3 ? 1 : 0;
  int left = n - tail, c = 1;
  while (left > 0) {
    const int take = std::min(std::min(c, chunk), left);
    sizes.push_back(take);
    left -= take;
    c *= 2;
  }
  if (tail) sizes.push_back(tail);
  return sizes;
}

// One chunk through the three-stage pipeline over two device slots: H2D (copy-in stream), kernels (compute stream) and
// D2H (copy-out stream) are chained by events, so H2D(k+1) and D2H(k-1) overlap compute(k) - across calls too, because
// the slot index and the events live in the ctx.  The CNN workspace is ordered by ctx->ev_ws inside cnn_forward.
int enqueue_chunk(vex_ctx* ctx, int mode, const uint8_t* in_host, uint8_t* out_host, int nf, int h, int w, int flags) {
  const size_t fin = (size_t)h * w * 3, fout = fin * 4;
  int rc = ensure_slots(ctx, fin * nf, fout * nf);
  if (rc) return rc;
  const unsigned long long k = ctx->chunk_seq;
  const int sl = (int)(k & 1);
  if (k >= 2) VEX_CUDA(ctx, cudaStreamWaitEvent(ctx->st_h2d, ctx->ev_comp[sl], 0));   // input slot consumed
  VEX_CUDA(ctx, cudaMemcpyAsync(ctx->slot_in[sl], in_host, fin * nf, cudaMemcpyHostToDevice, ctx->st_h2d));
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
  VEX_CUDA(ctx, cudaMemcpyAsync(out_host, ctx->slot_out[sl], fout * nf, cudaMemcpyDeviceToHost, ctx->st_d2h));
  VEX_CUDA(ctx, cudaEventRecord(ctx->ev_d2h[sl], ctx->st_d2h));
  ctx->chunk_seq = k + 1;
  return VEX_OK;
}

// Host staging of a chunk: pageable frames -> pinned buffer.  One thread copies ~5-10 GB/s, which for a 4- or 8-frame
// chunk (25-50 MB) is as long as the kernels of the previous chunk; up to 4 threads share the frames of a chunk.
void stage_frames(uint8_t* dst, const uint8_t* const* frames, int nf, size_t fin) {
  const int nt = std::min(nf, 4);
  if (nt <= 1) {
    for (int i = 0; i < nf; ++i) memcpy(dst + fin * i, frames[i], fin);
    return;
  }
  std::vector<std::thread> th;
  for (int t = 1; t < nt; ++t)
    th.emplace_back([=] { for (int i = t; i < nf; i += nt) memcpy(dst + fin * i, frames[i], fin); });
  for (int i = 0; i < nf; i += nt) memcpy(dst + fin * i, frames[i], fin);
  for (std::thread& x : th) x.join();
}

int drain(vex_ctx* ctx) {
  if (!ctx->st_h2d) return VEX_OK;
  VEX_CUDA(ctx, cudaStreamSynchronize(ctx->st_d2h));
  VEX_CUDA(ctx, cudaStreamSynchronize(ctx->st_comp));
  VEX_CUDA(ctx, cudaStreamSynchronize(ctx->st_h2d));
  return VEX_OK;
}
}  // namespace

extern "C" {

int vex_abi_version(void) { return VEX_ABI_VERSION; }

#ifndef VEX_SOURCE_HASH
#define VEX_SOURCE_HASH "unknown"
#endif
const char* vex_build_info(void) {
  return "vex-src:" VEX_SOURCE_HASH " sm_100a"
#ifdef VEX_DEBUG_PROBES
         " probes"
#endif
      ;
}

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
  {
    vex::DeviceGuard guard(device_ordinal);      // creates the primary context if needed, then restores the caller's device
    if (!guard.ok) {
      g_create_error = "vex_create: cudaSetDevice failed";
      return VEX_ERR_CUDA;
    }
    cudaFree(0);
  }
  vex_ctx* ctx = new vex_ctx();
  ctx->device = device_ordinal;
  ctx->sm_count = prop.multiProcessorCount;
  *out = ctx;
  return VEX_OK;
}

int vex_destroy(vex_ctx* ctx) {
  if (!ctx) return VEX_ERR_INVALID;
  {
    vex::DeviceGuard guard(ctx->device);
    cudaDeviceSynchronize();
    vex::cnn_destroy(ctx);
    vex::rrdb_destroy(ctx);
    if (ctx->ev_ws_rrdb) cudaEventDestroy(ctx->ev_ws_rrdb);
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
    for (cudaEvent_t e : ctx->ev_ticket)
      if (e) cudaEventDestroy(e);
    if (ctx->ev_ws) cudaEventDestroy(ctx->ev_ws);
    for (uint8_t* b : ctx->stage_in)
      if (b) cudaFreeHost(b);
  }
  delete ctx;
  return VEX_OK;
}

const char* vex_last_error(const vex_ctx* ctx) {
  return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

int vex_bicubic2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w,
                     int round_mode, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
  return vex::launch_bicubic2x(ctx, in_hwc, out_hwc, n, h, w, round_mode, (cudaStream_t)stream);
}

int vex_lanczos4_2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
  return vex::launch_lanczos4_2x(ctx, in_hwc, out_hwc, n, h, w, (cudaStream_t)stream);
}

int vex_resize2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, int interpolation,
                    void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
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
  VEX_ON_DEVICE(ctx);
  return vex::launch_bilateral5(ctx, in_hwc, out_hwc, n, h, w, sigma_color, sigma_space, (cudaStream_t)stream);
}

int vex_filter3x3_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, const float* kernel9,
                     void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
  return vex::launch_filter3x3(ctx, in_hwc, out_hwc, n, h, w, kernel9, (cudaStream_t)stream);
}

int vex_cnn_load_weights(vex_ctx* ctx, const vex_cnn_desc* arch, const float* const* tensors,
                         int n_tensors) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
  const int rc = vex::cnn_load_weights(ctx, arch, tensors, n_tensors);
  if (rc == VEX_OK) vex::cnn_noalloc_setup(ctx);    // resolves the static scratch of vex_cnn2x_u8_noalloc; cannot fail the load
  return rc;
}

int vex_cnn2x_u8_noalloc(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w,
                         int in_bgr, int out_bgr, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_hwc || n <= 0 || h <= 0 || w <= 0)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_cnn2x_u8_noalloc: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::cnn_forward_noalloc(ctx, in_hwc, out_hwc, n, h, w, in_bgr != 0, out_bgr != 0, (cudaStream_t)stream);
}

int vex_cnn2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w,
                 int in_bgr, int out_bgr, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_hwc || n <= 0 || h <= 0 || w <= 0)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_cnn2x_u8: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::cnn_forward(ctx, in_hwc, out_hwc, nullptr, 0, n, h, w, in_bgr != 0, out_bgr != 0,
                          (cudaStream_t)stream);
}

int vex_cnn2x_f32_debug(vex_ctx* ctx, const uint8_t* in_hwc, float* out_nchw, int n, int h, int w,
                        int in_bgr, int tap, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_nchw || n <= 0 || h <= 0 || w <= 0 || tap < 0 || tap > 6 + VEX_CNN_MAX_EXTRA)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_cnn2x_f32_debug: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::cnn_forward(ctx, in_hwc, nullptr, out_nchw, tap, n, h, w, in_bgr != 0, 0,
                          (cudaStream_t)stream);
}

int vex_cnn2x_f32in(vex_ctx* ctx, const float* in_nchw, uint8_t* out_hwc, float* out_nchw, int n, int h, int w,
                    int out_bgr, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_nchw || (!out_hwc && !out_nchw) || n <= 0 || h <= 0 || w <= 0)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_cnn2x_f32in: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::cnn_forward(ctx, reinterpret_cast<const uint8_t*>(in_nchw), out_hwc, out_nchw, 0, n, h, w,
                          vex::IN_F32_NCHW, out_bgr != 0, (cudaStream_t)stream);
}

int vex_bicubic2x_f32in(vex_ctx* ctx, const float* in_nchw, uint8_t* out_hwc, int n, int h, int w, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_nchw || !out_hwc || n <= 0 || h <= 0 || w <= 0)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_bicubic2x_f32in: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::launch_bicubic2x_f32in(ctx, in_nchw, out_hwc, n, h, w, (cudaStream_t)stream);
}

int vex_cnn_profile_stats(vex_ctx* ctx, float* mean7, float* min7, float* max7, int* n_forwards) {
  if (!ctx || !n_forwards) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
  return vex::cnn_profile_stats(ctx, mean7, min7, max7, n_forwards);
}

int vex_rrdb_load_weights(vex_ctx* ctx, const vex_rrdb_desc* arch, const float* const* tensors, int n_tensors) {
  if (!ctx) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
  return vex::rrdb_load_weights(ctx, arch, tensors, n_tensors);
}

int vex_rrdb_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, int in_bgr, int out_bgr,
                void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_hwc || n <= 0 || h <= 0 || w <= 0) return vex::fail(ctx, VEX_ERR_INVALID, "vex_rrdb_u8: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::rrdb_forward(ctx, in_hwc, out_hwc, nullptr, 0, n, h, w, in_bgr != 0, out_bgr != 0, (cudaStream_t)stream);
}

int vex_rrdb_frame_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int h, int w, int tile_size, int tile_pad,
                      int in_bgr, int out_bgr, void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_hwc || h <= 0 || w <= 0) return vex::fail(ctx, VEX_ERR_INVALID, "vex_rrdb_frame_u8: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::rrdb_frame(ctx, in_hwc, out_hwc, h, w, tile_size, tile_pad, in_bgr != 0, out_bgr != 0, (cudaStream_t)stream);
}

int vex_rrdb_f32_debug(vex_ctx* ctx, const uint8_t* in_hwc, float* out_nchw, int n, int h, int w, int in_bgr, int tap,
                       void* stream) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_hwc || !out_nchw || n <= 0 || h <= 0 || w <= 0 || tap < 0)
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_rrdb_f32_debug: bad argument");
  VEX_ON_DEVICE(ctx);
  return vex::rrdb_forward(ctx, in_hwc, nullptr, out_nchw, tap, n, h, w, in_bgr != 0, 0, (cudaStream_t)stream);
}

int vex_rrdb_workspace_bytes(const vex_ctx* ctx, int n, int h, int w, size_t* out) {
  if (!ctx || !out || n <= 0 || h <= 0 || w <= 0) return VEX_ERR_INVALID;
  *out = vex::rrdb_workspace_bytes(ctx, n, h, w);
  return VEX_OK;
}

int vex_workspace_bytes(const vex_ctx* ctx, int n, int h, int w, size_t* out) {
  if (!ctx || !out || n <= 0 || h <= 0 || w <= 0) return VEX_ERR_INVALID;
  *out = vex::cnn_workspace_bytes(ctx, n, h, w);
  return VEX_OK;
}

int vex_host_alloc(vex_ctx* ctx, size_t bytes, void** out) {
  if (!ctx || !out || bytes == 0) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
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
  VEX_ON_DEVICE(ctx);
  const size_t fin = (size_t)h * w * 3, fout = fin * 4;
  int f0 = 0;
  if (int rc = ensure_slots(ctx, fin * std::min(chunk, n), fout * std::min(chunk, n))) return rc;   // grow once, not per chunk
  for (int nf : chunk_schedule(n, chunk)) {
    int rc = enqueue_chunk(ctx, mode, in_host + fin * f0, out_host + fout * f0, nf, h, w, flags);
    if (rc) return rc;
    f0 += nf;
  }
  return drain(ctx);
}

int vex_submit(vex_ctx* ctx, int mode, const uint8_t* in_host, uint8_t* out_host, int n, int h, int w, int flags,
               vex_ticket* ticket) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!in_host || !out_host || !ticket || n <= 0 || h <= 0 || w <= 0 ||
      (mode != VEX_MODE_BICUBIC && mode != VEX_MODE_CNN))
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_submit: bad argument");
  VEX_ON_DEVICE(ctx);
  const unsigned long long t = ctx->next_ticket;
  const int ri = (int)(t % vex_ctx::kTickets);
  if (!ctx->ev_ticket[ri]) VEX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_ticket[ri], cudaEventDisableTiming));
  if (ctx->ticket_in_ring[ri]) {                 // the ring entry still belongs to an un-waited older ticket
    VEX_CUDA(ctx, cudaEventSynchronize(ctx->ev_ticket[ri]));
    ctx->ticket_in_ring[ri] = 0;
  }
  int rc = enqueue_chunk(ctx, mode, in_host, out_host, n, h, w, flags);
  if (rc) return rc;
  VEX_CUDA(ctx, cudaEventRecord(ctx->ev_ticket[ri], ctx->st_d2h));
  ctx->ticket_in_ring[ri] = t;
  ctx->next_ticket = t + 1;
  *ticket = t;
  return VEX_OK;
}

int vex_wait(vex_ctx* ctx, vex_ticket ticket) {
  if (!ctx) return VEX_ERR_INVALID;
  if (ticket == 0 || ticket >= ctx->next_ticket) return vex::fail(ctx, VEX_ERR_INVALID, "vex_wait: unknown ticket");
  VEX_ON_DEVICE(ctx);
  const int ri = (int)(ticket % vex_ctx::kTickets);
  if (ctx->ticket_in_ring[ri] != ticket) return VEX_OK;      // already waited for, or completed before its entry was reused
  VEX_CUDA(ctx, cudaEventSynchronize(ctx->ev_ticket[ri]));
  ctx->ticket_in_ring[ri] = 0;
  return VEX_OK;
}

int vex_upscale_frames(vex_ctx* ctx, int mode, const uint8_t* const* frames_in, uint8_t* out_host, int n, int h, int w,
                       int chunk, int flags) {
  if (!ctx) return VEX_ERR_INVALID;
  if (!frames_in || !out_host || n <= 0 || h <= 0 || w <= 0 || chunk <= 0 ||
      (mode != VEX_MODE_BICUBIC && mode != VEX_MODE_CNN))
    return vex::fail(ctx, VEX_ERR_INVALID, "vex_upscale_frames: bad argument");
  for (int i = 0; i < n; ++i)
    if (!frames_in[i]) return vex::fail(ctx, VEX_ERR_INVALID, "vex_upscale_frames: null frame");
  VEX_ON_DEVICE(ctx);
  const size_t fin = (size_t)h * w * 3, fout = fin * 4;
  const std::vector<int> sizes = chunk_schedule(n, chunk);
  const size_t need = fin * (size_t)*std::max_element(sizes.begin(), sizes.end());
  if (int rc = ensure_slots(ctx, need, need * 4)) return rc;
  if (need > ctx->stage_in_bytes) {
    int rc = drain(ctx);                         // nothing may still be reading the old staging buffers
    if (rc) return rc;
    for (int i = 0; i < 2; ++i) {
      if (ctx->stage_in[i]) cudaFreeHost(ctx->stage_in[i]);
      ctx->stage_in[i] = nullptr;
    }
    ctx->stage_in_bytes = 0;
    for (int i = 0; i < 2; ++i) VEX_CUDA(ctx, cudaHostAlloc(reinterpret_cast<void**>(&ctx->stage_in[i]), need, cudaHostAllocDefault));
    ctx->stage_in_bytes = need;
  }
  int f0 = 0;
  for (int nf : sizes) {
    const int sl = (int)(ctx->chunk_seq & 1);
    // the staging buffer of this slot was last read by the H2D of chunk seq-2; the host copy below runs while the GPU
    // works on chunk seq-1
    if (ctx->chunk_seq >= 2) VEX_CUDA(ctx, cudaEventSynchronize(ctx->ev_h2d[sl]));
    stage_frames(ctx->stage_in[sl], frames_in + f0, nf, fin);
    int rc = enqueue_chunk(ctx, mode, ctx->stage_in[sl], out_host + fout * f0, nf, h, w, flags);
    if (rc) return rc;
    f0 += nf;
  }
  return drain(ctx);
}

int vex_cnn_profile_ms(vex_ctx* ctx, float* ms7) {
  if (!ctx || !ms7) return VEX_ERR_INVALID;
  VEX_ON_DEVICE(ctx);
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
  else if (!strcmp(name, "nvtx")) slot = &ctx->opt_nvtx;
  else if (!strcmp(name, "conv1_variant")) slot = &ctx->opt_conv1_variant;
  else if (!strcmp(name, "pdl")) slot = &ctx->opt_pdl;
  else if (!strcmp(name, "rrdb_batch")) slot = &ctx->opt_rrdb_batch;
  else if (!strcmp(name, "rrdb_zigzag")) slot = &ctx->opt_rrdb_zigzag;
  else if (!strcmp(name, "rrdb_half")) slot = &ctx->opt_rrdb_half;
  else if (!strcmp(name, "rrdb_promo128")) slot = &ctx->opt_rrdb_promo128;
  else if (!strcmp(name, "rrdb_stages")) slot = &ctx->opt_rrdb_stages;
  if (!slot) return vex::fail(ctx, VEX_ERR_INVALID, std::string("vex_set_option: unknown ") + name);
#ifndef VEX_DEBUG_PROBES
  // timing probes that skip work and therefore produce wrong pixels exist only in -DVEX_DEBUG_PROBES builds
  if ((slot == &ctx->opt_thin_probe && value != 0) || (slot == &ctx->opt_desc_base_offset && value != 0) ||
      (slot == &ctx->opt_epi_variant && (value & 7) != 0))
    return vex::fail(ctx, VEX_ERR_UNSUPPORTED, std::string("vex_set_option: ") + name +
                     " is a wrong-result bring-up probe; rebuild with VEX_DEBUG_PROBES=1 to use it");
#endif
  const int prev = *slot;
  *slot = value;
  if (slot == &ctx->opt_conv3_bn) {
    VEX_ON_DEVICE(ctx);
    if (int rc = vex::cnn_repack(ctx)) return rc;
  }
  if (slot == &ctx->opt_profile && value) vex::cnn_profile_reset(ctx);
  return prev;
}

}  // extern "C"
