#include <cstdio>
#include <cstdint>
__global__ void k(uint32_t* o){
  int i0=0x11,i1=0x22,i2=0x33,i3=0x44; uint32_t lo,r, c=0xAABBCCDD, r2;
  asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, 0;":"=r"(lo):"r"(i1),"r"(i0));
  asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;":"=r"(r):"r"(i3),"r"(i2),"r"(lo));
  asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;":"=r"(r2):"r"(i3),"r"(i2),"r"(c));
  o[0]=lo; o[1]=r; o[2]=r2;
  float f0=17.9f,f1=-3.f,f2=300.f,f3=254.99f;
  int a0=__float2int_rz(f0),a1=__float2int_rz(f1),a2=__float2int_rz(f2),a3=__float2int_rz(f3);
  asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, 0;":"=r"(lo):"r"(a1),"r"(a0));
  asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;":"=r"(r):"r"(a3),"r"(a2),"r"(lo));
  o[3]=r;
}
int main(){uint32_t* d; cudaMalloc(&d,64); k<<<1,1>>>(d); uint32_t h[4]; cudaMemcpy(h,d,16,cudaMemcpyDeviceToHost); printf("lo=%08x r=%08x r2=%08x f=%08x\n",h[0],h[1],h[2],h[3]); return 0;}
