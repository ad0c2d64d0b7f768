// HBM write ceiling on B200: pure 16-byte stores with different cache operators, cudaMemset, and a 1:4 read:write mix.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE>
__global__ void wr(uint4* __restrict__ dst, const uint4* __restrict__ src, size_t n16) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
    uint4 v = make_uint4(1, 2, 3, (uint32_t)i);
    if (MODE == 4) v = __ldg(src + (i >> 2));            // 1 read : 4 written
    uint4* p = dst + i;
    if (MODE == 0 || MODE == 4) *p = v;
    if (MODE == 1) asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (MODE == 2) asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (MODE == 3) {   // 32-byte store (sm_100): thread i writes chunks 2i, 2i+1 of the first half, then the same in the second half
      if (i < n16 / 2) asm volatile("st.global.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%1,%2,%3,%4};" ::"l"(dst + 2 * i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
  }
}
template <int MODE>
void run(const char* name, uint4* dst, uint4* src, size_t bytes, int blocks) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0);
    wr<MODE><<<blocks, 256>>>(dst, src, bytes / 16);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
  }
  printf("%-28s blocks=%6d  %.3f ms  write %.0f GB/s%s\n", name, blocks, best, bytes / best / 1e6, MODE == 4 ? " (+25% read)" : "");
}
int main() {
  const size_t bytes = 300ull * 2160 * 3840 * 3;
  uint4 *dst, *src; cudaMalloc(&dst, bytes); cudaMalloc(&src, bytes / 4); cudaMemset(src, 1, bytes / 4);
  for (int blocks : {148 * 8, 148 * 32, 1 << 20}) {
    run<0>("st.v4 default", dst, src, bytes, blocks);
    run<1>("st.cs", dst, src, bytes, blocks);
    run<2>("st.L1::no_allocate", dst, src, bytes, blocks);
    run<3>("st.v8 32B evict_first (x2)", dst, src, bytes, blocks);
    run<4>("ld 1 : st 4", dst, src, bytes, blocks);
  }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int r = 0; r < 3; ++r) { cudaEventRecord(e0); cudaMemsetAsync(dst, 0, bytes); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); printf("cudaMemset %.3f ms %.0f GB/s\n", ms, bytes / ms / 1e6); }
  return 0;
}
