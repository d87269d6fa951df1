#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c){
  unsigned long long d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__global__ void k1(float* out, int iters){
  float a[8]; for(int i=0;i<8;i++) a[i]=threadIdx.x*0.001f+i;
  float m=0.999f, c=0.001f;
  for(int it=0; it<iters; it++){
#pragma unroll
    for(int i=0;i<8;i++) a[i]=fmaf(a[i],m,c);
  }
  float s=0; for(int i=0;i<8;i++) s+=a[i]; out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k2(float* out, int iters){
  unsigned long long a[8]; for(int i=0;i<8;i++){ float2 f=make_float2(threadIdx.x*0.001f+i, i*0.5f); a[i]=*reinterpret_cast<unsigned long long*>(&f);}
  float2 mf=make_float2(0.999f,0.999f), cf=make_float2(0.001f,0.001f);
  unsigned long long m=*reinterpret_cast<unsigned long long*>(&mf), c=*reinterpret_cast<unsigned long long*>(&cf);
  for(int it=0; it<iters; it++){
#pragma unroll
    for(int i=0;i<8;i++) a[i]=fma2(a[i],m,c);
  }
  float s=0; for(int i=0;i<8;i++){ float2 f=*reinterpret_cast<float2*>(&a[i]); s+=f.x+f.y;} out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
int main(){
  float* d; cudaMalloc(&d, 148*8*256*4);
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int iters=100000;
  for(int rep=0;rep<2;rep++){
    cudaEventRecord(e0); k1<<<148*8,256>>>(d,iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms,e0,e1);
    double inst=148.0*8*8*iters*8; // warp instrs
    printf("FFMA : %.3f ms  %.2f warp-inst/clk/SM (1.965GHz)  %.1f TFLOP/s\n", ms, inst/(ms*1e-3)/148/1.965e9, inst*64/(ms*1e-3)/1e12);
    cudaEventRecord(e0); k2<<<148*8,256>>>(d,iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms,e0,e1);
    printf("FFMA2: %.3f ms  %.2f warp-inst/clk/SM  %.1f TFLOP/s\n", ms, inst/(ms*1e-3)/148/1.965e9, inst*128/(ms*1e-3)/1e12);
  }
  return 0;
}
