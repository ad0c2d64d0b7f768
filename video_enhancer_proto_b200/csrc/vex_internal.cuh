// Internal declarations shared by the translation units of libvex.so (not part of the ABI).
#pragma once

#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/vex.h"

namespace vex {

struct CnnPlan;
struct RrdbPlan;
int launch_filter3x3(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, const float* k9, cudaStream_t s);
int launch_bilateral5(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, float sigma_color,
                      float sigma_space, cudaStream_t s);
int launch_lanczos4_2x(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, cudaStream_t s);
int launch_resize2x(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, int linear, cudaStream_t s);

// cnn.cu

}  // namespace vex

struct vex_ctx {
  int device = 0;
  float* bilateral_cw = nullptr;   // colour-weight table of vex_bilateral5_u8 (768 floats, device)
  float bilateral_sigma = -1.f;    // sigma_color the table was built for
  int sm_count = 148;
  std::string err;
  unsigned long long launches = 0;

  // streaming path (vex_upscale_host): H2D / compute / D2H streams over 2 device slots
  cudaStream_t st_h2d = nullptr, st_comp = nullptr, st_d2h = nullptr;
  cudaEvent_t ev_h2d[2] = {nullptr, nullptr}, ev_comp[2] = {nullptr, nullptr}, ev_d2h[2] = {nullptr, nullptr};
  uint8_t* slot_in[2] = {nullptr, nullptr};
  uint8_t* slot_out[2] = {nullptr, nullptr};
  size_t slot_in_bytes = 0, slot_out_bytes = 0;
  // asynchronous tickets (vex_submit / vex_wait): ticket t completes when ev_ticket[t % VEX_MAX_TICKETS] fires
  static constexpr int kTickets = 8;
  cudaEvent_t ev_ticket[kTickets] = {};
  unsigned long long next_ticket = 1;                 // tickets are 1-based; 0 is "none"
  unsigned long long ticket_in_ring[kTickets] = {};   // which ticket last used the ring entry (0 = never / waited)
  unsigned long long chunk_seq = 0;                   // chunks enqueued so far (slot = seq & 1)
  // pinned staging of vex_upscale_frames (pageable per-frame pointers in): one buffer per device slot
  uint8_t* stage_in[2] = {nullptr, nullptr};
  size_t stage_in_bytes = 0;
  // every CNN forward on this ctx shares one activation workspace: `ev_ws` marks the end of the last forward, whatever
  // stream it ran on, and the next forward waits for it (two streams must not interleave on the workspace)
  cudaEvent_t ev_ws = nullptr;
  bool ws_used = false;
  unsigned attr_mask3 = 0;        // same for the conv_tc3 instantiations
  unsigned attr_mask = 0;         // which kernels had their max-dynamic-smem attribute set on this device

  // CNN state
  vex::CnnPlan* cnn = nullptr;
  // Real-ESRGAN RRDBNet state (rrdb.cu)
  vex::RrdbPlan* rrdb = nullptr;
  cudaEvent_t ev_ws_rrdb = nullptr;
  bool ws_rrdb_used = false;

  // debug knobs (vex_set_option)
  int opt_desc_base_offset = 0;   // smem-descriptor base_offset field: 0 (measured correct on B200), 1: (addr>>7)&7
  int opt_force_generic_bicubic = 0;
  int opt_force_generic_resize = 0;  // vex_resize2x_u8: byte-indexed kernel even when the 4-pixel word kernel applies
  int opt_thin_probe = 0;            // im2col27 timing probes (wrong results), see thin_tc.cuh
  int opt_conv_variant = 0;          // 0 = conv_tc2 (cluster pairs, double-buffered TMEM), 1 = conv_tc v1
  int max_pairs3 = 0;                // same for conv_tc3
  int max_pairs = 0;                 // co-resident 2-CTA clusters of conv_tc2 (queried once)
  int opt_bicubic_rows = 0;          // rows per strip override (0 = auto)
  int opt_conv3_bn = 256;         // Cout block of conv3 on the tensor cores (128 or 256)
  int opt_fuse_tail = 4;          // upconv1 + upconv2 in one kernel: 4 = tail_tc3.cuh (default), 1 = tail_tc.cuh, 3 = tail_tc2.cuh; 0 = two kernels with `up` in HBM
  int opt_thin_tc = 1;            // thin layers on tcgen05 (thin_tc.cuh); 0 = fp32 CUDA-core kernels (cnn_small.cuh)
  int opt_epi_variant = 0;        // bring-up timing probes for the conv_tc epilogue (produce wrong results)
  int opt_conv_dbg = 0;           // bring-up: print per-role cycle counters of every conv_tc launch (synchronises)
  int opt_profile = 0;            // record per-layer events in cnn_forward (vex_cnn_profile_ms / vex_cnn_profile_stats)
  int opt_nvtx = 0;               // one NVTX range per layer launch
  int opt_rrdb_stages = 0;        // A/B: cap on the activation stages of the RRDB conv kernel (0 = as many as fit, up to 4)
  int opt_rrdb_half = 1;          // RRDB conv: the half-filled last chunk of a 96 / 160-channel input is loaded as 32 channels; 0 = A/B
  int opt_rrdb_promo128 = 1;      // RRDB conv layers that read part of a pixel's channels use 128-byte L2 promotion; 0 = A/B
  int opt_rrdb_zigzag = 1;        // RRDB conv launches alternate their sweep direction (L2 reuse between consecutive layers); 0 = A/B
  int opt_rrdb_batch = 16;        // Real-ESRGAN tile loop: most tiles of a frame that run as one batch of canvases (1 = one after the other)
  int opt_pdl = 1;                // RRDB layer chain: programmatic dependent launch (a layer's prologue overlaps its predecessor's tail)
  int opt_conv1_variant = 0;      // 0 = fp16 offset-1024 operand builder (MODE_CONV1_H), 1 = bf16 builder with I2F conversions
  int opt_ws_budget_mb = 49152;   // CNN activation workspace cap; larger batches run in sub-batches
};

namespace vex {

// Entry points run on ctx->device and leave the caller's current device as they found it (PyTorch reads it back).
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};
#define VEX_ON_DEVICE(ctx)                                                                   \
  ::vex::DeviceGuard guard__((ctx)->device);                                                 \
  if (!guard__.ok) return ::vex::fail((ctx), VEX_ERR_CUDA, "cudaSetDevice failed")

inline int fail(vex_ctx* ctx, int code, const std::string& msg) {
  if (ctx) ctx->err = msg;
  return code;
}

#define VEX_CUDA(ctx, call)                                                                  \
  do {                                                                                       \
    cudaError_t e__ = (call);                                                                \
    if (e__ != cudaSuccess) {                                                                \
      return ::vex::fail((ctx), VEX_ERR_CUDA,                                                \
                         std::string(#call) + ": " + cudaGetErrorString(e__) + " (" +        \
                             __FILE__ + ":" + std::to_string(__LINE__) + ")");               \
    }                                                                                        \
  } while (0)

// bicubic.cu
int launch_bicubic2x(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w,
                     int round_mode, cudaStream_t s);

int launch_bicubic2x_f32in(vex_ctx* ctx, const float* in_nchw, uint8_t* out, int n, int h, int w, cudaStream_t s);

// cnn.cu: the no-allocation entry (static device scratch)
void cnn_noalloc_setup(vex_ctx* ctx);
int cnn_forward_noalloc(vex_ctx* ctx, const uint8_t* in, uint8_t* out_u8, int n, int h, int w, int in_bgr, int out_bgr,
                        cudaStream_t s);
// rrdb.cu
int rrdb_load_weights(vex_ctx* ctx, const vex_rrdb_desc* arch, const float* const* tensors, int n);
int rrdb_forward(vex_ctx* ctx, const uint8_t* in, uint8_t* out_u8, float* out_f32, int tap, int n, int h, int w,
                 int in_bgr, int out_bgr, cudaStream_t s);
int rrdb_frame(vex_ctx* ctx, const uint8_t* in, uint8_t* out, int h, int w, int tile_size, int tile_pad, int in_bgr,
               int out_bgr, cudaStream_t s);
size_t rrdb_workspace_bytes(const vex_ctx* ctx, int n, int h, int w);
void rrdb_destroy(vex_ctx* ctx);

// cnn.cu
constexpr int IN_F32_NCHW = 2;   // cnn_forward's in_bgr argument: 0 RGB u8, 1 BGR u8, 2 = `in` is fp32 NCHW in [0,1] (no swap)
void cnn_profile_reset(vex_ctx* ctx);
int cnn_profile_stats(vex_ctx* ctx, float* mean7, float* min7, float* max7, int* n_forwards);
int cnn_load_weights(vex_ctx* ctx, const vex_cnn_desc* arch, const float* const* tensors, int n);
int cnn_forward(vex_ctx* ctx, const uint8_t* in, uint8_t* out_u8, float* out_f32, int tap,
                int n, int h, int w, int in_bgr, int out_bgr, cudaStream_t s);
size_t cnn_workspace_bytes(const vex_ctx* ctx, int n, int h, int w);
void cnn_destroy(vex_ctx* ctx);
int cnn_repack(vex_ctx* ctx);
int cnn_profile_ms(vex_ctx* ctx, float* ms7);   // after opt_conv3_bn changed

}  // namespace vex
