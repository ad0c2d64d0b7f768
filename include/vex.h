/*
 * vex.h — C ABI of libvex.so: the B200-native (sm_100a) per-frame 2x upscaler hot path.
 *
 * The reference (mindburnlabs/video-enhancer-proto) is pure Python and has NO FFI for this
 * path; its boundary is the duck-typed GPUVideoEnhancer / handler object.  This ABI is what
 * the Python mirror of that object (video_enhancer_proto_b200/upscaler.py) binds with ctypes.
 * Each entry point cites the reference code it replaces (paths relative to the reference repo).
 *
 * Conventions
 *   - every function returns 0 (VEX_OK) or a negative vex_status; nothing throws or exits;
 *   - vex_ctx is opaque, one per device, NOT thread-safe (the reference holds one global
 *     `enhancer`, app.py:624, and max_concurrent_tasks=1, agents/enhancer/video_enhancer_sota.py:72);
 *   - frame buffers are caller-owned, packed uint8 HWC, no row padding, frames contiguous;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream); device-pointer
 *     entry points are asynchronous on it; CNN calls on one ctx share its activation workspace and
 *     must therefore be stream-ordered with respect to each other;
 *   - there is NO CPU fallback: without a usable sm_100 device vex_create fails.
 */
#ifndef VEX_H_
#define VEX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VEX_ABI_VERSION 2

typedef enum vex_status {
  VEX_OK = 0,
  VEX_ERR_INVALID = -1,      /* bad argument (null pointer, non-positive size, bad flag) */
  VEX_ERR_CUDA = -2,         /* a CUDA runtime/driver call failed; see vex_last_error */
  VEX_ERR_NO_WEIGHTS = -3,   /* CNN entry point called before vex_cnn_load_weights */
  VEX_ERR_UNSUPPORTED = -4,  /* device is not sm_100, or architecture plan not supported */
  VEX_ERR_NOMEM = -5
} vex_status;

typedef struct vex_ctx vex_ctx;

/* ---- lifetime --------------------------------------------------------------------------- */

/* Replaces GPUVideoEnhancer.__init__ / load_models (app.py:245-263). */
int vex_create(int device_ordinal, vex_ctx** out);
int vex_destroy(vex_ctx* ctx);
/* Owned by ctx, valid until the next call on ctx.  ctx may be NULL (returns the create error). */
const char* vex_last_error(const vex_ctx* ctx);
int vex_abi_version(void);
/* "vex-src:<sha256 of the sources the library was compiled from> sm_100a[ probes]" (static string).  The Python binding
 * compares the hash with the sources it sits next to, so a stale prebuilt library cannot be loaded silently. */
const char* vex_build_info(void);

/* ---- bicubic fallback ------------------------------------------------------------------- */

/* 2x bicubic (Keys cubic, A=-0.75, align_corners=False, border replicate), per channel.
 * Replaces GPUVideoEnhancer._enhance_frame_cpu_fallback (app.py:359-385) for round_mode 0
 * (clip then TRUNCATE, app.py:376) and cv2.resize(..., INTER_CUBIC)
 * (models/enhancement/frame/basic_upscaler_fallback.py:142, realesrgan_fallback.py:104) for
 * round_mode 1 (round half up).  in: n*[h][w][3] u8 device; out: n*[2h][2w][3] u8 device. */
int vex_bicubic2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc,
                     int n, int h, int w, int round_mode, void* stream);

/* ---- the other per-frame stages of BasicUpscalerFallback --------------------
 * models/enhancement/frame/basic_upscaler_fallback.py:113-128 runs, per frame,
 *   LANCZOS mode:  cv2.bilateralFilter(frame, 5, 50, 50) -> cv2.resize(INTER_LANCZOS4) -> cv2.filter2D(sharpen) -> clip
 *   CUBIC mode:    cv2.resize(INTER_CUBIC) (= vex_bicubic2x_u8 round_mode 1)           -> cv2.filter2D(sharpen) -> clip
 * All three take and return packed u8 HWC device buffers, n frames, any channel order (they are channel-symmetric). */

/* cv2.resize(frame, (2w, 2h), interpolation=cv2.INTER_LANCZOS4) (basic_upscaler_fallback.py:52,120) for 8-bit images:
 * OpenCV's 11-bit fixed-point 8-tap separable path, replicate border.  Bit-exact.  out: n*[2h][2w][3]. */
int vex_lanczos4_2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, void* stream);

/* cv2.resize(frame, (2w, 2h), interpolation=code) for every method BasicUpscalerFallback maps a model name to
 * (basic_upscaler_fallback.py:51-57,120); `interpolation` takes OpenCV's own flag values: 0 INTER_NEAREST, 1 INTER_LINEAR
 * (11-bit fixed-point bilinear, replicate border), 2 INTER_CUBIC (= vex_bicubic2x_u8 round_mode 1), 3 INTER_AREA (enlarging
 * by 2 replicates pixels, like NEAREST), 4 INTER_LANCZOS4 (= vex_lanczos4_2x_u8).  NEAREST / LINEAR / AREA / LANCZOS4 are
 * bit-exact, CUBIC within 1 LSB.  out: n*[2h][2w][3]. */
int vex_resize2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, int interpolation,
                    void* stream);

/* cv2.bilateralFilter(frame, 5, sigma_color, sigma_space) (basic_upscaler_fallback.py:117; the reference passes 50, 50):
 * diameter 5 -> the 13 taps within radius 2, BORDER_REFLECT_101.  in != out.  out: n*[h][w][3]. */
int vex_bilateral5_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, float sigma_color,
                      float sigma_space, void* stream);

/* cv2.filter2D(img, -1, kernel) for a 3x3 fp32 kernel given row-major (basic_upscaler_fallback.py:123-127: the reference's
 * kernel is 0.1 * [[-1,-1,-1],[-1,9,-1],[-1,-1,-1]]), BORDER_REFLECT_101, rounded and saturated to u8 (which is also what
 * the reference's np.clip(...).astype(np.uint8) leaves).  in != out.  out: n*[h][w][3]. */
int vex_filter3x3_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w, const float* kernel9,
                     void* stream);

/* ---- CNN upscaler ------------------------------------------------------------------------ */

/* Channel plan of the AdvancedUpscaler family (app.py:269-301):
 *   conv1 3->c1, conv2 c1->c2 (+skip, needs c1==c2), conv3 c2->c3, conv4 c3->c4, conv5 c4->3,
 *   upconv1 ConvT 3->up_ch k4 s2 p1, upconv2 ConvT up_ch->3 k3 s1 p1.
 * The reference is {128,128,256,128,64}.  Widths must be multiples of 128 in [128, 512]; up_ch 64.
 * n_extra (0 for the reference, at most VEX_CNN_MAX_EXTRA) deepens the plan for BASELINE.json configs[4]:
 * after conv3, n_extra more blocks conv3x3 c3->c3 + BatchNorm + ReLU of the same kind as app.py:293. */
typedef struct vex_cnn_desc {
  int c1, c2, c3, c4, up_ch;
  float bn_eps;              /* 1e-5 (nn.BatchNorm2d default, app.py:284-287) */
  int n_extra;
} vex_cnn_desc;

#define VEX_CNN_NUM_TENSORS 30
#define VEX_CNN_MAX_EXTRA 4
#define VEX_CNN_TENSORS_PER_EXTRA 6
/* tensors[i]: fp32 host arrays in the reference module's state_dict order (app.py:273-287):
 *   conv1..conv5 {weight,bias}, upconv1 {weight,bias}, upconv2 {weight,bias},
 *   bn1..bn4 {weight,bias,running_mean,running_var}     (num_batches_tracked is skipped),
 *   then per extra block k = 1..n_extra: conv3x<k> {weight,bias}, bn3x<k> {weight,bias,running_mean,running_var};
 *   n_tensors = VEX_CNN_NUM_TENSORS + VEX_CNN_TENSORS_PER_EXTRA * n_extra.
 * Folds eval-mode BN into per-channel scale/shift, converts weights to bf16 in the kernels'
 * layouts, uploads.  Copies; the caller keeps ownership.  Replaces _create_gpu_upscaler
 * (app.py:265-311) minus the random init (the caller supplies the values). */
int vex_cnn_load_weights(vex_ctx* ctx, const vex_cnn_desc* arch,
                         const float* const* tensors, int n_tensors);

/* n frames u8 HWC [h][w][3] -> u8 HWC [2h][2w][3], device pointers.
 * in_bgr/out_bgr: channel order of the buffers relative to the network's RGB planes
 * (the batch path takes RGB and returns BGR, app.py:469,550; the single-frame path does not
 * swap, app.py:334-346).  Replaces _process_frame_batch_gpu (app.py:529-553) and
 * _enhance_frame_with_gpu (app.py:322-352) including clip(x*255,0,255).astype(uint8). */
int vex_cnn2x_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc,
                 int n, int h, int w, int in_bgr, int out_bgr, void* stream);

/* vex_cnn2x_u8 for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels and the same output bytes as vex_cnn2x_u8; the activation workspace is a static device
 * array inside the library instead of a cudaMalloc'd one, and the call makes no cudaMalloc / cudaFree / event creation.
 *   limit:    one frame's workspace (vex_workspace_bytes(ctx, 1, h, w)) must fit VEX_NOALLOC_SCRATCH_BYTES = 176 MiB,
 *             i.e. up to 480 x 360 (BASELINE configs[0]) with the reference widths (1040 B per input pixel); larger
 *             frames return VEX_ERR_NOMEM with the two byte counts in vex_last_error.  n frames run in sub-batches
 *             of as many as fit.  Every GPU whose context loads the library holds the 176 MiB until process exit.
 *   set-up:   vex_cnn_load_weights (which allocates) must have run on this ctx; it also resolves the scratch address.
 *             Options "profile" and "conv_dbg" allocate and are refused here (VEX_ERR_UNSUPPORTED).
 *   sharing:  ONE array per device, shared by every call and every ctx of the process, holding the last call's
 *             contents (nothing needs clearing: each layer writes its output before it is read).  Calls are ordered
 *             only by `stream`; two calls that overlap on the GPU race on the scratch.  It is independent of the
 *             workspace of vex_cnn2x_u8, so a no-alloc call and an ordinary call may overlap. */
#define VEX_NOALLOC_SCRATCH_BYTES (176u << 20)
int vex_cnn2x_u8_noalloc(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc,
                         int n, int h, int w, int in_bgr, int out_bgr, void* stream);

/* fp32 input form of vex_cnn2x_u8 for the tensor branch of the single-frame entry (app.py:337-338: a caller may hand
 * `_enhance_frame_with_gpu` an fp32 NCHW tensor already in [0,1], which the reference feeds to conv1 unquantised).
 * in_nchw: n*[3][h][w] fp32 device, any real values; conv1 takes each value as a hi + lo bf16 pair (relative input error
 * 2^-17), the weights are the same bf16 images as the uint8 path minus the 1/255 fold.  No colour swap on input (the
 * reference does none on this branch).  out_hwc (u8, n*[2h][2w][3]) and out_nchw (fp32 pre-quantisation clamp output,
 * n*[3][2h][2w]) may each be NULL, not both. */
int vex_cnn2x_f32in(vex_ctx* ctx, const float* in_nchw, uint8_t* out_hwc, float* out_nchw,
                    int n, int h, int w, int out_bgr, void* stream);

/* fp32 input form of vex_bicubic2x_u8 (round_mode 0) for the tensor branch of _enhance_frame_cpu_fallback
 * (app.py:367-376): in_nchw n*[3][h][w] fp32 in [0,1] (any real values) -> clip(bicubic*255, 0, 255) truncated,
 * out_hwc n*[2h][2w][3] u8.  fp32 arithmetic in ATen's order (x pass inside each row, then y). */
int vex_bicubic2x_f32in(vex_ctx* ctx, const float* in_nchw, uint8_t* out_hwc, int n, int h, int w, void* stream);

/* Parity hook: same network, pre-quantisation fp32 output, NCHW [n][3][2h][2w] (device).
 * `tap` selects what is written: 0 = final clamp(upconv2) (app.py:299); 1..5 = x1..x5
 * (NCHW fp32, [n][C][h][w]; with n_extra > 0, x3 is the output of conv3 itself); 6 = relu(upconv1)
 * ([n][up_ch][2h][2w]); 7 + k = output of extra block k + 1 ([n][c3][h][w]). */
int vex_cnn2x_f32_debug(vex_ctx* ctx, const uint8_t* in_hwc, float* out_nchw,
                        int n, int h, int w, int in_bgr, int tap, void* stream);

/* Per-layer device time (ms, CUDA events on the caller's stream) of the most recent vex_cnn2x_u8 /
 * vex_cnn2x_f32_debug(tap 0) call made while option "profile" is 1:
 * ms7 = {conv1, conv2, conv3 (+ extra blocks), conv4, conv5, upconv1, upconv2}; sub-batches are summed.  Synchronises on
 * the last event.  (The reference has no per-layer timing; its PerformanceTracker,
 * utils/performance_monitor.py:178, only wraps whole videos.) */
int vex_cnn_profile_ms(vex_ctx* ctx, float* ms7);

/* Per-layer device time over EVERY forward made since option "profile" was last set to 1 (bench.py's roofline uses the
 * mean over the whole timed region, not one step): mean7 / min7 / max7 in ms per forward (sub-batches summed), same
 * layer order as vex_cnn_profile_ms; n_forwards = forwards folded in.  Any of the three arrays may be NULL.
 * Synchronises on the last recorded event. */
int vex_cnn_profile_stats(vex_ctx* ctx, float* mean7, float* min7, float* max7, int* n_forwards);

/* Bytes of device workspace the ctx holds (grows on demand) for a batch of this shape. */
int vex_workspace_bytes(const vex_ctx* ctx, int n, int h, int w, size_t* out);

/* ---- Real-ESRGAN RRDBNet ------------------------------------------------- */

/* RealESRGANNetwork (models/enhancement/vsr/realesrgan_handler.py:59-100) as the handler builds it (:143-150):
 * num_feat 64, num_grow_ch 32 fixed; num_block 23 and scale 4 in the reference, scale 2 and other depths accepted. */
typedef struct vex_rrdb_desc {
  int num_feat, num_grow_ch, num_block, scale;
} vex_rrdb_desc;

/* tensors[i]: fp32 host arrays in the reference module's state_dict order: conv_first {weight,bias},
 * body.<b>.rdb<1..3>.conv<1..5> {weight,bias} for b = 0..num_block-1, conv_body, conv_up1, (conv_up2 for scale 4),
 * conv_hr, conv_last; n_tensors = 2 + 30*num_block + 2 + 2*(scale == 4 ? 2 : 1) + 4.  Converts to fp16 in the kernels'
 * layouts (conv_first: bf16 hi + lo with the /255 of :311 folded in) and uploads; the caller keeps ownership.
 * Replaces RealESRGANHandler._ensure_model_loaded / _load_model (:134-212) minus the download. */
int vex_rrdb_load_weights(vex_ctx* ctx, const vex_rrdb_desc* arch, const float* const* tensors, int n_tensors);

/* n images u8 HWC [h][w][3] -> u8 HWC [s*h][s*w][3] (s = scale), device pointers, the whole image in one pass:
 * RealESRGANHandler._process_single (:308-322) incl. /255, clamp(0,1), x255 and the TRUNCATING uint8 cast.
 * in_bgr / out_bgr: channel order of the buffers relative to the network's RGB planes (the handler converts BGR->RGB
 * before and RGB->BGR after, :290,300; pass 1,1 for cv2 frames).  fp16 operands and activations, fp32 accumulation. */
int vex_rrdb_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int n, int h, int w,
                int in_bgr, int out_bgr, void* stream);

/* One frame the way RealESRGANHandler._enhance_frame does it (:286-306): if h or w exceeds tile_size the frame is cut
 * into tile_size tiles with tile_pad pixels of context, each tile runs through the network on its own and its un-padded
 * core is pasted into the output (_tile_process, :324-372; the reference uses tile_size 512, tile_pad 10) - cropping and
 * pasting are strided device copies; otherwise the whole frame is one pass.  Device pointers, packed u8 HWC. */
int vex_rrdb_frame_u8(vex_ctx* ctx, const uint8_t* in_hwc, uint8_t* out_hwc, int h, int w, int tile_size, int tile_pad,
                      int in_bgr, int out_bgr, void* stream);

/* Parity hook: fp32 NCHW of the selected activation.  tap 0 = clamp(conv_last, 0, 1) [n][3][s*h][s*w];
 * 1 = conv_first [n][64][h][w]; 2 = feat + conv_body(body(feat)) [n][64][h][w]; 3 = lrelu(conv_hr) [n][64][s*h][s*w];
 * 10 + b = output of body block b [n][64][h][w]; 1001..1004 = x1..x4 of body[0].rdb1 [n][32][h][w], 1005 = its output. */
int vex_rrdb_f32_debug(vex_ctx* ctx, const uint8_t* in_hwc, float* out_nchw, int n, int h, int w, int in_bgr, int tap,
                       void* stream);

/* Bytes of activation workspace a batch of this shape needs (grows on demand). */
int vex_rrdb_workspace_bytes(const vex_ctx* ctx, int n, int h, int w, size_t* out);

/* ---- host-buffer streaming entry (the end-to-end path) ------------------------------------ */

#define VEX_MODE_BICUBIC 0
#define VEX_MODE_CNN 1

/* Pinned host memory for frame pools (cudaHostAlloc); frames decoded into it stream without
 * an extra staging copy. */
int vex_host_alloc(vex_ctx* ctx, size_t bytes, void** out);
int vex_host_free(vex_ctx* ctx, void* p);

/* n frames from HOST memory to HOST memory: chunks of `chunk` frames are copied H2D, upscaled
 * and copied D2H on double-buffered streams so copy and compute overlap; returns when every
 * output frame is in out_host.  Pageable pointers work (slower); vex_host_alloc'd ones overlap.
 * flags: bit0 round_mode (bicubic) / in_bgr (cnn); bit1 out_bgr (cnn).
 * Replaces the frame loop body of _process_video_gpu_decorated (app.py:458-482) between
 * cap.read() and out.write(). */
int vex_upscale_host(vex_ctx* ctx, int mode, const uint8_t* in_host, uint8_t* out_host,
                     int n, int h, int w, int chunk, int flags);

/* Asynchronous form of vex_upscale_host for callers that produce frames incrementally (the decoder loop of
 * app.py:458-482): vex_submit enqueues H2D -> kernels -> D2H of ONE chunk of n frames on the ctx's copy-in / compute /
 * copy-out streams (device slot = submit count & 1, chained by events exactly like the chunks of vex_upscale_host) and
 * returns at once with a ticket; vex_wait blocks until that chunk's output is in out_host.  Chunks complete in submit
 * order.  in_host must stay untouched until the chunk's H2D has run and out_host until vex_wait returns; pinned
 * (vex_host_alloc) buffers make both copies asynchronous.  Up to 8 tickets may be outstanding; a 9th submit first
 * waits for the oldest.  Waiting twice, or for a ticket older than the last 8, returns VEX_OK. */
typedef unsigned long long vex_ticket;
int vex_submit(vex_ctx* ctx, int mode, const uint8_t* in_host, uint8_t* out_host,
               int n, int h, int w, int flags, vex_ticket* ticket);
int vex_wait(vex_ctx* ctx, vex_ticket ticket);

/* The plugin-call form (`_process_frame_batch_gpu(frames_batch)`, app.py:529-553: a Python list of separately
 * allocated, pageable frames in): frames_in[i] points at frame i (packed u8 HWC [h][w][3], host, pageable or pinned).
 * Frames are staged chunk by chunk into the ctx's own pinned buffers while the previous chunk computes (chunk sizes
 * ramp 1, 2, 4, ... up to `chunk`, last chunk 1 frame, so the exposed staging / first H2D / last D2H are one frame
 * each), outputs go to out_host (n*[2h][2w][3]; pinned memory overlaps).  Returns when every output is in out_host. */
int vex_upscale_frames(vex_ctx* ctx, int mode, const uint8_t* const* frames_in, uint8_t* out_host,
                       int n, int h, int w, int chunk, int flags);

/* Counters for bench.py's `gpu_launches`: kernels launched by this ctx since creation. */
int vex_kernel_launches(const vex_ctx* ctx, unsigned long long* out);

/* Knobs.  Returns the previous value, or a negative vex_status.
 *   product:   "profile" (0/1: per-layer events; setting 1 clears the statistics of vex_cnn_profile_stats), "nvtx" (0/1:
 *              one NVTX range per layer launch), "ws_budget_mb" (activation workspace cap; larger
 *              batches run as sub-batches), "conv3_bn" (128/256: Cout block of conv3), "pdl" (0/1: programmatic dependent launch of the
 *              Real-ESRGAN layer chain), "rrdb_batch" (how many tiles of a Real-ESRGAN frame run as one batch of canvases;
 *              1 = one after the other as in realesrgan_handler.py:342-370; bit-identical for every value).
 *   variants kept as bit-identity cross-checks (tests/test_cnn_gpu.py): "conv_variant" (0 = default v3 cta_group::2,
 *              1 = single-CTA, 2 = multicast pairs), "conv1_variant" (0 = fp16 offset-1024 operand builder, 1 = bf16 builder), "thin_tc" (0 = fp32 CUDA-core thin layers), "fuse_tail" (default 4 = fused tail with upconv2's vertical
 *              tap accumulated by the tensor core; 1 = the first fused tail (27 planes + 9-tap gather; same products, other fp32
 *              summation order, within 1 LSB of 4), 3 = 1 with `up` kept in tensor memory (bit-identical to 1), 0 = upconv1 / upconv2 as two
 *              kernels with `up` in HBM), "force_generic_bicubic", "force_generic_resize", "bicubic_rows".
 *   bring-up probes that PRODUCE WRONG RESULTS (timing experiments only; refused unless the library was built with
 *              -DVEX_DEBUG_PROBES): "epi_variant", "thin_probe", "desc_base_offset"; "conv_dbg" prints per-role cycle
 *              counters and synchronises. */
int vex_set_option(vex_ctx* ctx, const char* name, int value);

#ifdef __cplusplus
}
#endif
#endif /* VEX_H_ */
