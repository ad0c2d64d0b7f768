// placeholder until the tcgen05 path lands (replaced in the next commit)
#include "vex_internal.cuh"
namespace vex {
int cnn_load_weights(vex_ctx* ctx, const vex_cnn_desc*, const float* const*, int) {
  return fail(ctx, VEX_ERR_UNSUPPORTED, "cnn path not built yet");
}
int cnn_forward(vex_ctx* ctx, const uint8_t*, uint8_t*, float*, int, int, int, int, int, int, cudaStream_t) {
  return fail(ctx, VEX_ERR_NO_WEIGHTS, "vex_cnn_load_weights has not been called");
}
size_t cnn_workspace_bytes(const vex_ctx*, int, int, int) { return 0; }
void cnn_destroy(vex_ctx*) {}
}  // namespace vex
