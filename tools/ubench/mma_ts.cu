// tcgen05.mma with the A operand in TENSOR MEMORY (D[tmem] = A[tmem] x B[smem]): which TMEM cell holds A[row][k]?
// A is written with tcgen05.st.32x32b (thread i of warp w -> lane 32w + i, register j -> column j) as bf16 pairs,
// B is a 0/1 selection matrix, so D[r][n] = A[r][n] exactly and the read-back shows the layout the tensor core assumes.
// Hypothesis tested: lane = row, column j of the operand's base = elements k = 2j (low half) and 2j + 1 (high half);
// a K = 16 step spans 8 columns, the next step starts 8 columns further.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../video_enhancer_proto_b200/csrc -o mma_ts mma_ts.cu && ./mma_ts
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace vex;

__device__ __forceinline__ void mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
               ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st_32x32_x16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}

// A[r][k] = r + k / 64 for k < 32 (two K = 16 steps); exactly representable in bf16 for r < 128? r + k/64 needs 7 + 6 bits: no.
// Use small integers instead: A[r][k] = (r % 16) * 2 + (k % 2) + 32 * (k / 2 % 4) ... keep it simple: value = code(r, k) < 256.
__host__ __device__ inline float code(int r, int k) { return (float)(((r * 7 + k * 3) % 251) - 125); }

__global__ void __launch_bounds__(128, 1) k(float* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (sbase - raw);
  __shared__ uint32_t slot;
  __shared__ __align__(8) unsigned long long bar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, r = threadIdx.x;
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::fence_mbar_init(); }
  // B: 32 rows (n) x 64 bf16 (k) K-major, SWIZZLE_128B image; B[n][k] = (n == k) for k < 32
  for (int i = threadIdx.x; i < 32 * 64; i += 128) {
    const int n = i / 64, kk = i % 64;
    const uint16_t one = 0x3F80;
    reinterpret_cast<uint16_t*>(sm)[n * 64 + (((kk >> 3) ^ (n & 7)) << 3) + (kk & 7)] = (n == kk) ? one : 0;
  }
  ptx::fence_proxy_async_smem();
  if (warp == 0) { ptx::tmem_alloc(ptx::smem_u32(&slot), 128); ptx::tmem_relinquish(); }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  // A in TMEM columns [0, 16): 32 bf16 per row
  uint32_t v[16];
  for (int j = 0; j < 16; ++j) v[j] = ptx::pack_bf16x2(code(r, 2 * j), code(r, 2 * j + 1));
  tmem_st_32x32_x16(tmem + ((uint32_t)(warp * 32) << 16), v);
  ptx::tmem_st_wait();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  if (warp == 1 && ptx::elect_one()) {
    const uint32_t idesc = ptx::idesc_bf16_f32(128, 32);
    for (int ks = 0; ks < 2; ++ks) {
      const uint64_t bd = ptx::smem_desc_sw128(sbase + ks * 32, 0);
      mma_bf16_ts(tmem + 32, tmem + ks * 8, bd, idesc, ks);
    }
    ptx::tc_commit(ptx::smem_u32(&bar));
  }
  ptx::mbar_wait(ptx::smem_u32(&bar), 0);
  ptx::tc_fence_after();
  uint32_t d[32];
  ptx::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + 32, d);
  ptx::tmem_ld_wait();
  for (int n = 0; n < 32; ++n) out[r * 32 + n] = __uint_as_float(d[n]);
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tmem, 128);
}

int main() {
  float* d; cudaMalloc(&d, 128 * 32 * 4);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192);
  k<<<1, 128, 8192>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
  float h[128 * 32];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int r = 0; r < 128; ++r)
    for (int n = 0; n < 32; ++n)
      if (h[r * 32 + n] != code(r, n)) { if (bad < 8) printf("D[%d][%d] = %g, expected A[%d][%d] = %g\n", r, n, h[r * 32 + n], r, n, code(r, n)); ++bad; }
  printf("mma_ts: %d mismatches of 4096 (0 = lane is the row, column j holds k = 2j | 2j+1, a K=16 step is 8 columns)\n", bad);
  return 0;
}
