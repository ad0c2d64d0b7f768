// tcgen05.mma issue-to-retire rate on B200: cta_group::1 (M=128) vs cta_group::2 (M=256 over a CTA pair), N=128/256,
// SS operands (K-major SWIZZLE_128B, resident in shared memory), back-to-back MMAs into 2 accumulators.
#include <cstdio>
#include <cuda_bf16.h>
#include "../../video_enhancer_proto_b200/csrc/ptx.cuh"
using namespace vex;
template <int CG, int N, int AOFF>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) k(long long* out, int reps) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw), sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  const uint32_t rank = ptx::cluster_ctarank();
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (32768 + 32768 + 8192) / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u;
  ptx::fence_proxy_async_smem();
  const uint32_t A = sbase, B = sbase + 40960, bar = sbase + 73728 + 8192, slot = bar + 8;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_mbar_init(); }
  if (warp == 0) { if (CG == 2) { ptx::tmem_alloc2(slot, 512); ptx::tmem_relinquish2(); } else { ptx::tmem_alloc(slot, 512); ptx::tmem_relinquish(); } }
  ptx::tc_fence_before(); __syncthreads(); ptx::cluster_sync(); ptx::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(sm + 73728 + 8192 + 8);
  long long t0 = 0, t1 = 0;
  if (threadIdx.x == 0 && (CG == 1 || rank == 0)) {
    const uint32_t idesc = ptx::idesc_bf16_f32(CG == 2 ? 256 : 128, N);
    t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t ad = ptx::smem_desc_sw128(A + AOFF * 128 + kk * 32, 0), bd = ptx::smem_desc_sw128(B + kk * 32, 0);
        if (CG == 2) ptx::mma2_bf16_ss(tmem + (r & 1) * N, ad, bd, idesc, 1u);
        else ptx::mma_bf16_ss(tmem + (r & 1) * N, ad, bd, idesc, 1u);
      }
    }
    if (CG == 2) ptx::tc_commit2_mc(bar, (uint16_t)3); else ptx::tc_commit(bar);
  }
  ptx::mbar_wait(bar, 0);
  t1 = clock64();
  if (threadIdx.x == 0 && (CG == 1 || rank == 0)) out[blockIdx.x] = t1 - t0;
  ptx::tc_fence_before(); __syncthreads(); ptx::cluster_sync();
  if (warp == 0) { if (CG == 2) ptx::tmem_dealloc2(tmem, 512); else ptx::tmem_dealloc(tmem, 512); }
}
template <int CG, int N, int AOFF>
void run(const char* name) {
  long long* d; cudaMalloc(&d, 148 * 8); cudaMemset(d, 0, 148 * 8);
  const int reps = 2000, smem = 100 * 1024;
  cudaFuncSetAttribute(k<CG, N, AOFF>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k<CG, N, AOFF><<<148, 128, smem>>>(d, reps);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, d, 148 * 8, cudaMemcpyDeviceToHost);
  double mx = 0; for (int i = 0; i < 148; ++i) if (h[i] > mx) mx = (double)h[i];
  const double per = mx / (reps * 4.0), flop_per_sm = 2.0 * 128 * N * 16;
  printf("%-34s %s  %.1f cycles per MMA (K=16)  -> %.0f flop/clk/SM\n", name, cudaGetErrorString(e), per, flop_per_sm / per);
  cudaFree(d);
}
int main() {
  run<1, 32, 0>("cta_group::1 M128 N32");
  run<1, 32, 33>("cta_group::1 M128 N32 A+33 rows");
  run<1, 64, 0>("cta_group::1 M128 N64");
  run<1, 64, 33>("cta_group::1 M128 N64 A+33 rows");
  run<1, 96, 0>("cta_group::1 M128 N96");
  run<1, 128, 0>("cta_group::1 M128 N128");
  run<1, 256, 0>("cta_group::1 M128 N256");
  run<1, 128, 33>("cta_group::1 M128 N128 A+33 rows");
  run<2, 128, 0>("cta_group::2 M256 N128");
  run<2, 256, 0>("cta_group::2 M256 N256");
  run<2, 128, 33>("cta_group::2 M256 N128 A+33 rows");
  return 0;
}
