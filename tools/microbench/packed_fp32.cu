// microbenchmark: scalar FFMA/FMUL/FADD vs packed FFMA2/FMUL2/FADD2 issue throughput on sm_100a
#include <cuda_runtime.h>
#include <cstdio>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a,float b){u64 r; asm("mov.b64 %0,{%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(u64 v,float&a,float&b){asm("mov.b64 {%0,%1},%2;":"=f"(a),"=f"(b):"l"(v));}
template<int MODE> __global__ void k(float* o,int iters,float s){
  float a[8]; u64 p[8];
  for(int i=0;i<8;i++){a[i]=threadIdx.x*0.001f+i; p[i]=pk(a[i],a[i]+1.f);}
  u64 S=pk(s,s), C=pk(0.5f,0.25f);
  for(int it=0;it<iters;it++){
#pragma unroll
    for(int i=0;i<8;i++){
      if(MODE==0) a[i]=fmaf(a[i],s,0.5f);
      if(MODE==1) asm volatile("fma.rn.f32x2 %0,%0,%1,%2;":"+l"(p[i]):"l"(S),"l"(C));
      if(MODE==2) {asm volatile("mul.rn.f32 %0,%0,%1;":"+f"(a[i]):"f"(s));}
      if(MODE==3) asm volatile("mul.rn.f32x2 %0,%0,%1;":"+l"(p[i]):"l"(S));
      if(MODE==4) {asm volatile("add.rn.f32 %0,%0,%1;":"+f"(a[i]):"f"(s));}
      if(MODE==5) asm volatile("add.rn.f32x2 %0,%0,%1;":"+l"(p[i]):"l"(S));
      if(MODE==6) {asm volatile("mul.rn.f32 %0,%0,%1;":"+f"(a[i]):"f"(s)); asm volatile("min.f32 %0,%0,%1;":"+f"(a[(i+4)&7]):"f"(s));}
    }
  }
  float r=0; for(int i=0;i<8;i++){float x,y; upk(p[i],x,y); r+=a[i]+x+y;}
  o[blockIdx.x*blockDim.x+threadIdx.x]=r;
}
template<int MODE> void run(const char*name,int per){
  float*o; cudaMalloc(&o,148*8*256*4);
  int iters=20000; cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148*8,256>>>(o,100,1.0001f); cudaDeviceSynchronize();
  cudaEventRecord(e0); k<MODE><<<148*8,256>>>(o,iters,1.0001f); cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms,e0,e1);
  double winst=(double)148*8*8*iters*8*per; // warp-instructions
  printf("%-10s %.3f ms  %.1f Gwarp-inst/s  (%.2f inst/clk/SM at 1.965GHz)\n",name,ms,winst/ms/1e6,winst/(ms*1e-3)/148/1.965e9);
  cudaFree(o);
}
int main(){ run<0>("FFMA",1); run<1>("FFMA2",1); run<2>("FMUL",1); run<3>("FMUL2",1); run<4>("FADD",1); run<5>("FADD2",1); run<6>("FMUL+FMNMX",2); return 0; }
