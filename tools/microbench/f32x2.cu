#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long pk(float a, float b){ unsigned long long r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(unsigned long long v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b){ unsigned long long r; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b){ unsigned long long r; asm("add.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}

template<int MODE> __global__ void k(float* out, int iters, float s){
  float a0=threadIdx.x*1e-3f, a1=a0+1, a2=a0+2, a3=a0+3, a4=a0+4,a5=a0+5,a6=a0+6,a7=a0+7;
  if (MODE==0){ // scalar mul+add, 8 independent chains
    for(int i=0;i<iters;i++){
      a0=__fadd_rn(a0,__fmul_rn(a0,s)); a1=__fadd_rn(a1,__fmul_rn(a1,s)); a2=__fadd_rn(a2,__fmul_rn(a2,s)); a3=__fadd_rn(a3,__fmul_rn(a3,s));
      a4=__fadd_rn(a4,__fmul_rn(a4,s)); a5=__fadd_rn(a5,__fmul_rn(a5,s)); a6=__fadd_rn(a6,__fmul_rn(a6,s)); a7=__fadd_rn(a7,__fmul_rn(a7,s));
    }
  } else if (MODE==1){ // packed
    unsigned long long p0=pk(a0,a1),p1=pk(a2,a3),p2=pk(a4,a5),p3=pk(a6,a7), ss=pk(s,s);
    for(int i=0;i<iters;i++){
      p0=add2(p0,mul2(p0,ss)); p1=add2(p1,mul2(p1,ss)); p2=add2(p2,mul2(p2,ss)); p3=add2(p3,mul2(p3,ss));
    }
    upk(p0,a0,a1);upk(p1,a2,a3);upk(p2,a4,a5);upk(p3,a6,a7);
  } else { // ffma scalar
    for(int i=0;i<iters;i++){
      a0=fmaf(a0,s,a0);a1=fmaf(a1,s,a1);a2=fmaf(a2,s,a2);a3=fmaf(a3,s,a3);a4=fmaf(a4,s,a4);a5=fmaf(a5,s,a5);a6=fmaf(a6,s,a6);a7=fmaf(a7,s,a7);
    }
  }
  out[blockIdx.x*blockDim.x+threadIdx.x]=a0+a1+a2+a3+a4+a5+a6+a7;
}
template<int MODE> void run(const char* name){
  float* d; cudaMalloc(&d, 148*8*256*4);
  int iters=20000; cudaEvent_t e0,e1; cudaEventCreate(&e0);cudaEventCreate(&e1);
  k<MODE><<<148*8,256>>>(d,100,1e-9f); cudaDeviceSynchronize();
  cudaEventRecord(e0); k<MODE><<<148*8,256>>>(d,iters,1e-9f); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms,e0,e1);
  double lane_ops = 148.0*8*256*iters*8*2; // mul+add per element
  printf("%s: %.3f ms, %.2f T lane-ops/s (mul+add counted 2)\n", name, ms, lane_ops/ms/1e9);
}
int main(){ run<0>("scalar fmul+fadd"); run<1>("packed f32x2 mul+add"); run<2>("scalar ffma"); return 0; }
