// Does ptxas keep mul.rn.f32x2 + add.rn.f32x2 unfused? a*b+c with a=b=1+2^-12, c=-(1+2^-11):
// separate roundings give 0, a fused multiply-add gives 2^-24.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long pk(float a, float b){ unsigned long long r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(unsigned long long v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b){ unsigned long long r; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b){ unsigned long long r; asm("add.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__global__ void k(const float* in, float* out){
  float a=in[0], b=in[1], c=in[2];
  out[0]=__fadd_rn(__fmul_rn(a,b),c);
  unsigned long long r=add2(mul2(pk(a,a),pk(b,b)),pk(c,c));
  float x,y; upk(r,x,y); out[1]=x; out[2]=y;
  out[3]=fmaf(a,b,c);
}
int main(){ float h[3]={1.0f+1.0f/4096, 1.0f+1.0f/4096, -(1.0f+1.0f/2048)}, o[4]; float *di,*dout; cudaMalloc(&di,12);cudaMalloc(&dout,16);
 cudaMemcpy(di,h,12,cudaMemcpyHostToDevice); k<<<1,1>>>(di,dout); cudaMemcpy(o,dout,16,cudaMemcpyDeviceToHost);
 printf("scalar unfused=%g packed=(%g,%g) fma=%g  (unfused expects 0, fused 5.96e-08)\n",o[0],o[1],o[2],o[3]); return 0; }
