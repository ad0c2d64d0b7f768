// Micro-benchmark: per-SM issue rate of the instructions the bicubic kernel is built from (B200, sm_100a).
// Prints lane-ops per clock per SM for each.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 4096
template <int OP>
__global__ void __launch_bounds__(1024) k(uint32_t* out, long long* cyc, uint32_t seed) {
  uint32_t r[8];
  uint64_t q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { r[i] = seed + threadIdx.x * 8 + i; q[i] = ((uint64_t)r[i] << 32) | (r[i] ^ 0x3f800000u); }
  uint32_t c = seed | 0x3f800001u;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) asm volatile("fma.rn.f32 %0, %0, 0f3F800001, %1;" : "+r"(r[i]) : "r"(c));
      if (OP == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(q[i]) : "l"(q[(i + 1) & 7]));
      if (OP == 2) { uint64_t w = 0x3F8000013F800001ull; asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(q[i]) : "l"(w), "l"(q[(i+1)&7])); }
      if (OP == 3) asm volatile("cvt.rzi.s32.f32 %0, %0;" : "+r"(r[i]));
      if (OP == 4) asm volatile("cvt.rn.f32.u32 %0, %0;" : "+r"(r[i]));
      if (OP == 5) asm volatile("prmt.b32 %0, %0, %1, 0x7250;" : "+r"(r[i]) : "r"(c));
      if (OP == 6) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %0;" : "+r"(r[i]) : "r"(c));
      if (OP == 7) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+r"(r[i]) : "r"(r[(i + 1) & 7]));
      if (OP == 8) asm volatile("mad.lo.s32 %0, %0, 67, %1;" : "+r"(r[i]) : "r"(c));
      if (OP == 9) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(q[i]) : "l"(q[(i + 1) & 7]));
      if (OP == 10) asm volatile("cvt.rzi.sat.u8.f32 %0, %0;" : "+r"(r[i]));
      if (OP == 11) asm volatile("min.f32 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
      if (OP == 12) asm volatile("lop3.b32 %0, %0, %1, %1, 0x96;" : "+r"(r[i]) : "r"(c));
      if (OP == 14) { int a0=__float2int_rz(__uint_as_float(r[i])), a1=__float2int_rz(__uint_as_float(r[(i+1)&7])); asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a1), "r"(a0)); }
      if (OP == 13) asm volatile("shf.r.wrap.b32 %0, %0, %1, 8;" : "+r"(r[i]) : "r"(c));
    }
  }
  long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc ^= r[i] ^ (uint32_t)q[i] ^ (uint32_t)(q[i] >> 32);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP>
void run(const char* name, int lanes_per_op) {
  uint32_t* out; long long* cyc;
  int nb = 148;
  cudaMalloc(&out, nb * 1024 * 4); cudaMalloc(&cyc, nb * 8);
  k<OP><<<nb, 1024>>>(out, cyc, 1); k<OP><<<nb, 1024>>>(out, cyc, 1);
  cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, cyc, nb * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < nb; ++i) avg += h[i]; avg /= nb;
  double ops = 1024.0 * ITERS * 8;
  printf("%-28s %7.1f thread-instr/clk/SM  (%7.1f scalar ops/clk/SM)  err=%d\n", name, ops / avg, ops * lanes_per_op / avg, (int)cudaGetLastError());
  cudaFree(out); cudaFree(cyc);
}
int main() {
  run<0>("FFMA imm", 1); run<7>("FFMA 3reg", 1); run<1>("FFMA2 3reg", 2); run<2>("FFMA2 const-pair", 2);
  run<9>("FADD2", 2); run<8>("IMAD imm", 1);
  run<3>("F2I.S32.TRUNC", 1); run<10>("F2I.U8.SAT", 1); run<4>("I2F.U32", 1); run<5>("PRMT", 1); run<6>("I2IP (cvt.pack)", 1);
  run<11>("FMNMX", 1); run<12>("LOP3", 1); run<13>("SHF", 1); run<14>("F2IP.U8.F32 (2 floats)", 2);
  return 0;
}
