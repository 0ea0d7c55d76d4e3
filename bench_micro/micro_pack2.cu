#include <cuda_runtime.h>
#include <cstdio>
// throughput of scalar vs packed fp32: each thread runs a dependent-free unrolled stream of ops
template<int PACK> __global__ void k(float* out, float a, float b, int iters){
  float2 v[8]; 
  for(int i=0;i<8;i++) v[i]=make_float2(threadIdx.x*0.001f+i, threadIdx.x*0.002f+i);
  float2 A=make_float2(a,a), B=make_float2(b,b);
  for(int it=0; it<iters; ++it){
    #pragma unroll
    for(int i=0;i<8;i++){
      if(PACK){ v[i]=__fmul2_rn(v[i],A); v[i]=__fadd2_rn(v[i],B); }
      else { v[i].x=__fmul_rn(v[i].x,a); v[i].y=__fmul_rn(v[i].y,a); v[i].x=__fadd_rn(v[i].x,b); v[i].y=__fadd_rn(v[i].y,b);}
    }
  }
  float s=0; for(int i=0;i<8;i++) s+=v[i].x+v[i].y;
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
int main(){
  float* out; cudaMalloc(&out, 148*8*256*4);
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for(int pack=0;pack<2;pack++){
    for(int rep=0;rep<2;rep++){
      cudaEventRecord(e0);
      if(pack) k<1><<<148*8,256>>>(out,1.0001f,0.0001f,4096); else k<0><<<148*8,256>>>(out,1.0001f,0.0001f,4096);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms,e0,e1);
      double flops = 148.0*8*256*4096*8*4; // fp32 ops (mul+add on 2 comps)
      printf("pack=%d  %.3f ms  %.2f Tfp32op/s\n",pack,ms,flops/ms/1e9);
    }
  }
  return 0;
}
