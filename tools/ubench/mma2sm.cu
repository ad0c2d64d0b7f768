// Probe: tcgen05.mma.cta_group::2 conventions on B200.  D[256][128] = A[256][64] * B[128][64]^T, bf16 -> fp32.
// CTA r of the pair holds A rows [128r, 128r+128) and B rows (N) [64r, 64r+64) at the same shared offsets;
// the leader (rank 0) issues the MMAs; each CTA reads its 128 TMEM lanes.  Prints max |D - ref|.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_bf16.h>
#include "../../video_enhancer_proto_b200/csrc/ptx.cuh"
using namespace vex;
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
k(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw), sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  const uint32_t rank = ptx::cluster_ctarank();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // A tile at 0 (16 KB), B half at 16384 (8 KB), barrier at 24576, tmem slot at 24584
  for (int i = threadIdx.x; i < 128 * 64; i += 128) {
    const int r = i / 64, kk = i % 64;
    *reinterpret_cast<__nv_bfloat16*>(sm + r * 128 + (((kk >> 3) ^ (r & 7)) << 4) + (kk & 7) * 2) = A[(rank * 128 + r) * 64 + kk];
  }
  for (int i = threadIdx.x; i < 64 * 64; i += 128) {
    const int r = i / 64, kk = i % 64;
    *reinterpret_cast<__nv_bfloat16*>(sm + 16384 + r * 128 + (((kk >> 3) ^ (r & 7)) << 4) + (kk & 7) * 2) = B[(rank * 64 + r) * 64 + kk];
  }
  ptx::fence_proxy_async_smem();
  const uint32_t bar = sbase + 24576, slot = sbase + 24584;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_mbar_init(); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(128) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(sm + 24584);
  if (rank == 0 && threadIdx.x == 0) {
    const uint32_t idesc = ptx::idesc_bf16_f32(256, 128);
    for (int kk = 0; kk < 4; ++kk) {
      const uint64_t ad = ptx::smem_desc_sw128(sbase + kk * 32, 0), bd = ptx::smem_desc_sw128(sbase + 16384 + kk * 32, 0);
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                   "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                   ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(kk) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
  }
  ptx::mbar_wait(bar, 0);
  ptx::tc_fence_after();
  for (int g = 0; g < 4; ++g) {
    uint32_t v[32];
    ptx::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + g * 32, v);
    ptx::tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(size_t)(rank * 128 + warp * 32 + lane) * 128 + g * 32 + j] = __uint_as_float(v[j]);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128) : "memory");
}
int main() {
  std::vector<__nv_bfloat16> hA(256 * 64), hB(128 * 64);
  std::vector<float> fA(256 * 64), fB(128 * 64), ref(256 * 128), hD(256 * 128);
  srand(1);
  for (size_t i = 0; i < hA.size(); ++i) { hA[i] = __float2bfloat16((rand() % 17 - 8) / 8.f); fA[i] = __bfloat162float(hA[i]); }
  for (size_t i = 0; i < hB.size(); ++i) { hB[i] = __float2bfloat16((rand() % 13 - 6) / 4.f); fB[i] = __bfloat162float(hB[i]); }
  for (int m = 0; m < 256; ++m) for (int n = 0; n < 128; ++n) { float s = 0; for (int kk = 0; kk < 64; ++kk) s += fA[m * 64 + kk] * fB[n * 64 + kk]; ref[m * 128 + n] = s; }
  __nv_bfloat16 *dA, *dB; float* dD;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, hD.size() * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xff, hD.size() * 4);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  k<<<2, 128, 32768>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  printf("launch: %s\n", cudaGetErrorString(e));
  cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
  double worst = 0; int bad = 0;
  for (int i = 0; i < 256 * 128; ++i) { double d = fabs(hD[i] - ref[i]); if (!(d <= 1e-3)) { if (bad < 5) printf("  D[%d][%d]=%g ref %g\n", i / 128, i % 128, hD[i], ref[i]); ++bad; } if (d > worst) worst = d; }
  printf("max err %g, mismatches %d of %d\n", worst, bad, 256 * 128);
  return 0;
}
