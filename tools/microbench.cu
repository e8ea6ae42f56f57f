// tools/microbench.cu -- B200 primitive rates that size the push+deposit design:
// fp64 FMA peak, global RED.ADD.F64 (L2) rate for several address patterns, and
// shared-memory fp64 atomic / load rates.  Not part of the product.
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s %s\n",#x,cudaGetErrorString(e)); return 1;}}while(0)

__global__ void k_dfma(double* out, int iters) {
  double a0=threadIdx.x*1e-3, a1=a0+1, a2=a0+2, a3=a0+3, a4=a0+4, a5=a0+5, a6=a0+6, a7=a0+7;
  const double b=1.0000001, c=1e-9;
  for (int i=0;i<iters;++i){ a0=a0*b+c; a1=a1*b+c; a2=a2*b+c; a3=a3*b+c; a4=a4*b+c; a5=a5*b+c; a6=a6*b+c; a7=a7*b+c; }
  out[blockIdx.x*blockDim.x+threadIdx.x]=a0+a1+a2+a3+a4+a5+a6+a7;
}
// pattern 0: every lane its own address (random over `span` doubles)
// pattern 1: groups of 16 lanes share an address, consecutive groups adjacent (sorted particles, 16 ppc)
// pattern 2: whole warp one address
// pattern 3: lanes consecutive addresses (coalesced)
__global__ void k_red(double* J, long long span, int pattern, int iters) {
  const unsigned tid = blockIdx.x*blockDim.x+threadIdx.x;
  unsigned long long s = tid*2654435761u+12345u;
  for (int i=0;i<iters;++i){
    s = s*6364136223846793005ULL+1442695040888963407ULL;
    long long a;
    if (pattern==0) a = (long long)((s>>20)%span);
    else if (pattern==1) a = (long long)((((unsigned long long)(tid>>4)*27+ (i%27)))%span);
    else if (pattern==2) a = (long long)((((unsigned long long)(tid>>5)*64+(i&63)))%span);
    else a = (long long)(((unsigned long long)tid + (unsigned long long)i*1315423911u)%span);
    atomicAdd(J+a, 1.0);
  }
}
__global__ void k_satom(double* out, int iters, int pattern) {
  __shared__ double sm[4096];
  for (int i=threadIdx.x;i<4096;i+=blockDim.x) sm[i]=0;
  __syncthreads();
  unsigned s = threadIdx.x*2654435761u+1;
  for (int i=0;i<iters;++i){
    s = s*1664525u+1013904223u;
    int a = pattern==0 ? (s>>8)&4095 : ((threadIdx.x>>4)*27 + (i%27))&4095;
    atomicAdd(&sm[a], 1.0);
  }
  __syncthreads();
  if (threadIdx.x==0) out[blockIdx.x]=sm[0]+sm[17];
}
__global__ void k_sint(unsigned long long* out, int iters) {
  __shared__ unsigned long long sm[4096];
  for (int i=threadIdx.x;i<4096;i+=blockDim.x) sm[i]=0;
  __syncthreads();
  unsigned s = threadIdx.x*2654435761u+1;
  for (int i=0;i<iters;++i){ s = s*1664525u+1013904223u; atomicAdd(&sm[(s>>8)&4095], 1ULL); }
  __syncthreads();
  if (threadIdx.x==0) out[blockIdx.x]=sm[0]+sm[17];
}
__global__ void k_lds(double* out, int iters, int pattern) {
  __shared__ double sm[4096];
  for (int i=threadIdx.x;i<4096;i+=blockDim.x) sm[i]=i;
  __syncthreads();
  double acc=0; int base = pattern==0 ? threadIdx.x : (threadIdx.x>>4);
  for (int i=0;i<iters;++i){
#pragma unroll
    for (int k=0;k<27;++k) acc += sm[(base + k*37 + i)&4095];
  }
  out[blockIdx.x*blockDim.x+threadIdx.x]=acc;
}
// DMMA m8n8k4 fp64 tensor-core MMA: 8 independent accumulator tiles per warp
__global__ void k_dmma(double* out, int iters) {
  double c[8][2];
#pragma unroll
  for (int t=0;t<8;++t){ c[t][0]=threadIdx.x*1e-3+t; c[t][1]=1.0+t; }
  double a = 1.0 + threadIdx.x*1e-9, b = 1e-9*threadIdx.x;
  for (int i=0;i<iters;++i){
#pragma unroll
    for (int t=0;t<8;++t)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[t][0]), "+d"(c[t][1]) : "d"(a), "d"(b));
  }
  double s=0;
#pragma unroll
  for (int t=0;t<8;++t) s+=c[t][0]+c[t][1];
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
// mixed: DFMA stream and DMMA stream in the same warp (do they share the fp64 pipe?)
__global__ void k_mixed(double* out, int iters) {
  double c[4][2];
#pragma unroll
  for (int t=0;t<4;++t){ c[t][0]=threadIdx.x*1e-3+t; c[t][1]=1.0+t; }
  double a = 1.0 + threadIdx.x*1e-9, b = 1e-9*threadIdx.x;
  double a0=threadIdx.x*1e-3, a1=a0+1, a2=a0+2, a3=a0+3, a4=a0+4, a5=a0+5, a6=a0+6, a7=a0+7;
  const double bb=1.0000001, cc=1e-9;
  for (int i=0;i<iters;++i){
#pragma unroll
    for (int t=0;t<4;++t)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[t][0]), "+d"(c[t][1]) : "d"(a), "d"(b));
    a0=a0*bb+cc; a1=a1*bb+cc; a2=a2*bb+cc; a3=a3*bb+cc; a4=a4*bb+cc; a5=a5*bb+cc; a6=a6*bb+cc; a7=a7*bb+cc;
    a0=a0*bb+cc; a1=a1*bb+cc; a2=a2*bb+cc; a3=a3*bb+cc; a4=a4*bb+cc; a5=a5*bb+cc; a6=a6*bb+cc; a7=a7*bb+cc;
  }
  double s=a0+a1+a2+a3+a4+a5+a6+a7;
#pragma unroll
  for (int t=0;t<4;++t) s+=c[t][0]+c[t][1];
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
// 64-bit warp shuffles
__global__ void k_shfl(double* out, int iters) {
  double v0=threadIdx.x, v1=v0+1, v2=v0+2, v3=v0+3;
  for (int i=0;i<iters;++i){
    v0 += __shfl_xor_sync(0xffffffffu, v0, 1); v1 += __shfl_xor_sync(0xffffffffu, v1, 2);
    v2 += __shfl_xor_sync(0xffffffffu, v2, 4); v3 += __shfl_xor_sync(0xffffffffu, v3, 8);
  }
  out[blockIdx.x*blockDim.x+threadIdx.x]=v0+v1+v2+v3;
}
int main(){
  cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr,0));
  printf("device %s SMs %d\n", pr.name, pr.multiProcessorCount);
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1); float ms;
  double* out; CK(cudaMalloc(&out, 1<<26));
  const int nb = pr.multiProcessorCount*8;
  { int it=20000; k_dfma<<<nb,256>>>(out,100); cudaEventRecord(e0); k_dfma<<<nb,256>>>(out,it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
    printf("DFMA: %.2f TFLOP/s fp64 (%.1f DFMA/clk/SM @1.9GHz-equiv)\n", 2.0*8*it*(double)nb*256/ms/1e9, 8.0*it*nb*256/(ms*1e-3)/pr.multiProcessorCount/1.9e9); }
  { int it=20000; k_dmma<<<nb,256>>>(out,100); cudaEventRecord(e0); k_dmma<<<nb,256>>>(out,it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
    const double nm = 8.0*it*(double)nb*8;   // warp-level MMAs
    printf("DMMA m8n8k4: %.2f TFLOP/s fp64 (%.2f clk/MMA/SM @1.9GHz-equiv)\n", 2.0*256*nm/ms/1e9, (ms*1e-3)*1.9e9*pr.multiProcessorCount/nm); }
  { int it=20000; k_mixed<<<nb,256>>>(out,100); cudaEventRecord(e0); k_mixed<<<nb,256>>>(out,it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
    printf("mixed 4 DMMA + 16 DFMA per iter: %.3f ms; DFMA-only equiv would be %.2f TFLOP/s, DMMA part %.2f TFLOP/s\n", ms, 2.0*16*it*(double)nb*256/ms/1e9, 2.0*256*4.0*it*nb*8/ms/1e9); }
  { int it=20000; k_shfl<<<nb,256>>>(out,100); cudaEventRecord(e0); k_shfl<<<nb,256>>>(out,it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
    printf("SHFL.64+DADD: %.1f G lane-shuffles(64b)/s (%.2f warp-shfl64/clk/SM @1.9GHz)\n", 4.0*it*(double)nb*256/ms/1e6, 4.0*it*nb*8/(ms*1e-3)/pr.multiProcessorCount/1.9e9); }
  double* J; const long long big = 1LL<<27; CK(cudaMalloc(&J, big*8)); CK(cudaMemset(J,0,big*8));
  const char* pn[4]={"distinct-random","16-lane-groups","warp-uniform","coalesced"};
  for (int span_i=0; span_i<2; ++span_i) {
    long long span = span_i==0 ? (1LL<<22) : big;   // 32 MB (L2 resident) / 1 GB
    for (int p=0;p<4;++p){ int it=200; k_red<<<nb,256>>>(J,span,p,10); cudaEventRecord(e0); k_red<<<nb,256>>>(J,span,p,it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
      printf("RED.F64 global span=%lldMB %-16s: %.1f G atomics/s\n", span*8>>20, pn[p], (double)it*nb*256/ms/1e6); }
  }
  for (int p=0;p<2;++p){ int it=2000; k_satom<<<nb,256>>>(out,10,p); cudaEventRecord(e0); k_satom<<<nb,256>>>(out,it,p); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
    printf("shared atomicAdd f64 %-16s: %.1f G atomics/s (%.2f /clk/SM @1.9GHz)\n", p==0?"random":"16-lane-groups", (double)it*nb*256/ms/1e6, (double)it*nb*256/(ms*1e-3)/pr.multiProcessorCount/1.9e9); }
  { int it=2000; k_sint<<<nb,256>>>((unsigned long long*)out,10); cudaEventRecord(e0); k_sint<<<nb,256>>>((unsigned long long*)out,it); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
    printf("shared atomicAdd u64 random          : %.1f G atomics/s\n", (double)it*nb*256/ms/1e6); }
  for (int p=0;p<2;++p){ int it=500; k_lds<<<nb,256>>>(out,10,p); cudaEventRecord(e0); k_lds<<<nb,256>>>(out,it,p); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms,e0,e1);
    printf("LDS.64 %-16s: %.1f G loads/s (%.1f B/clk/SM @1.9GHz)\n", p==0?"lane-consecutive":"16-lane-bcast", 27.0*it*nb*256/ms/1e6, 27.0*8*it*nb*256/(ms*1e-3)/pr.multiProcessorCount/1.9e9); }
  return 0;
}
