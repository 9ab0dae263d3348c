// Micro-probe 2: how the FP32 min (ALU pipe) shares issue/dispatch with scalar FFMA and packed
// FFMA2 on B200 (sm_100a).  Decides between the packed and the scalar form of the 1-NN inner loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/issue_probe tools/issue_probe.cu
// Output unit: cycles per "2 pairs" body per warp scheduler = elapsed * clock / (bodies per quadrant).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){u64 r; asm("mov.b64 %0,{%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ u64 fma2(u64 a,u64 b,u64 c){u64 r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;":"=l"(r):"l"(a),"l"(b),"l"(c)); return r;}
__device__ __forceinline__ u64 add2(u64 a,u64 b){u64 r; asm volatile("add.rn.f32x2 %0,%1,%2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ float sfma(float a,float b,float c){float r; asm volatile("fma.rn.f32 %0,%1,%2,%3;":"=f"(r):"f"(a),"f"(b),"f"(c)); return r;}
__device__ __forceinline__ float sadd(float a,float b){float r; asm volatile("add.rn.f32 %0,%1,%2;":"=f"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ float smin3(float a,float b,float c){float r; asm volatile("min.f32 %0,%1,%2,%3;":"=f"(r):"f"(a),"f"(b),"f"(c)); return r;}
__device__ __forceinline__ float smin2(float a,float b){float r; asm volatile("min.f32 %0,%1,%2;":"=f"(r):"f"(a),"f"(b)); return r;}
constexpr int NACC = 8, ITERS = 2048;
// Every mode's body stands for "2 query-reference pairs" of the 1-NN loop.
template <int MODE> __global__ void __launch_bounds__(256) probe(float* out, float s, const float* sm_in) {
  __shared__ float sh[256];
  sh[threadIdx.x] = sm_in ? sm_in[threadIdx.x] : s;
  __syncthreads();
  u64 acc[NACC]; float a0[NACC], a1[NACC], mn[NACC];
  for (int i=0;i<NACC;i++){acc[i]=pk(threadIdx.x*1e-3f+i, i*0.5f); a0[i]=i+threadIdx.x; a1[i]=i*2.f; mn[i]=1e30f;}
  u64 b = pk(s, s*1.0001f), c = pk(s*0.5f, s*0.25f);
  float fb = s, fc = s*0.5f, fd = s*0.25f;
  int ia = threadIdx.x;
  for (int it=0; it<ITERS; ++it) {
#pragma unroll
    for (int i=0;i<NACC;i++) {
      if (MODE==0 || MODE==1 || MODE==2 || MODE==3 || MODE==4 || MODE==10 || MODE==11) {  // packed: 3 FFMA2
        acc[i]=fma2(b,b,acc[i]); acc[i]=fma2(c,c,acc[i]); acc[i]=fma2(b,c,acc[i]);
        float lo,hi; asm("mov.b64 {%0,%1},%2;":"=f"(lo),"=f"(hi):"l"(acc[i]));
        if (MODE==1) mn[i]=smin3(mn[i],lo,hi);
        if (MODE==2) { mn[i]=smin2(mn[i],lo); }                         // one 2-input min only
        if (MODE==3) { asm volatile("add.s32 %0,%0,%1;":"+r"(ia):"r"(__float_as_int(lo))); }  // one IADD
        if (MODE==4) { float v; asm volatile("ld.shared.f32 %0,[%1];":"=f"(v):"r"((unsigned)__cvta_generic_to_shared(sh + ((it+i)&255)))); mn[i]+=v; }
        if (MODE==10) { mn[i]=smin3(mn[i],lo,hi); mn[i]=smin3(mn[i],hi,lo); }  // two FMNMX3
        if (MODE==11) { asm volatile("{.reg .pred p; setp.lt.f32 p,%1,%0; @p mov.f32 %0,%1;}":"+f"(mn[i]):"f"(lo)); }
      }
      if (MODE==5 || MODE==6 || MODE==7 || MODE==8 || MODE==9) {  // scalar: 6 FFMA
        a0[i]=sfma(fb,fb,a0[i]); a1[i]=sfma(fb,fc,a1[i]);
        a0[i]=sfma(fc,fc,a0[i]); a1[i]=sfma(fc,fd,a1[i]);
        a0[i]=sfma(fb,fc,a0[i]); a1[i]=sfma(fd,fd,a1[i]);
        if (MODE==6) mn[i]=smin3(mn[i],a0[i],a1[i]);
        if (MODE==7) { mn[i]=smin2(mn[i],a0[i]); mn[i]=smin2(mn[i],a1[i]); }
        if (MODE==8) { asm volatile("add.s32 %0,%0,%1;":"+r"(ia):"r"(__float_as_int(a0[i]))); }
        if (MODE==9) { mn[i]=smin3(mn[i],a0[i],a1[i]); mn[i]=smin3(mn[i],a1[i],a0[i]); }
      }
      if (MODE==12) { mn[i]=smin3(mn[i],fb,fc); fb+=0.f; }             // FMNMX3 alone (plus nothing)
      if (MODE==13) {                                                   // exact-mode packed body: 3 FADD2 + FMUL2 + 2 FFMA2 + FMNMX3
        u64 dx=add2(acc[i],b), dy=add2(acc[i],c), dz=add2(b,acc[i]);
        u64 d; asm volatile("mul.rn.f32x2 %0,%1,%1;":"=l"(d):"l"(dx)); d=fma2(dy,dy,d); d=fma2(dz,dz,d);
        float lo,hi; asm("mov.b64 {%0,%1},%2;":"=f"(lo),"=f"(hi):"l"(d));
        mn[i]=smin3(mn[i],lo,hi); acc[i]=d;
      }
      if (MODE==14) {                                                   // exact-mode scalar body
        float dx0=sadd(a0[i],fb), dy0=sadd(a0[i],fc), dz0=sadd(a0[i],fd);
        float dx1=sadd(a1[i],fb), dy1=sadd(a1[i],fc), dz1=sadd(a1[i],fd);
        float d0; asm volatile("mul.rn.f32 %0,%1,%1;":"=f"(d0):"f"(dx0)); d0=sfma(dy0,dy0,d0); d0=sfma(dz0,dz0,d0);
        float d1; asm volatile("mul.rn.f32 %0,%1,%1;":"=f"(d1):"f"(dx1)); d1=sfma(dy1,dy1,d1); d1=sfma(dz1,dz1,d1);
        mn[i]=smin3(mn[i],d0,d1); a0[i]=d0; a1[i]=d1;
      }
    }
  }
  float r=ia; for(int i=0;i<NACC;i++){float x,y; asm("mov.b64 {%0,%1},%2;":"=f"(x),"=f"(y):"l"(acc[i])); r+=x+y+a0[i]+a1[i]+mn[i];}
  out[blockIdx.x*blockDim.x+threadIdx.x]=r;
}
template <int MODE> void run(const char* name, int clock_khz) {
  float* out; cudaMalloc(&out, 148*8*256*sizeof(float));
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  probe<MODE><<<148*8,256>>>(out,1.0001f,nullptr); cudaDeviceSynchronize();
  float best=1e9f;
  for (int rep=0; rep<3; ++rep) {
    cudaEventRecord(a); probe<MODE><<<148*8,256>>>(out,1.0001f,nullptr); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms,a,b); if (ms<best) best=ms;
  }
  // bodies per warp scheduler: 8 CTAs x 8 warps / 4 quadrants = 16 warps, each ITERS*NACC bodies
  double bodies = 16.0*ITERS*NACC;
  double cyc = best*1e-3*clock_khz*1e3/bodies;
  printf("%-44s %.3f ms  %.2f cycles per 2-pair body (at %d MHz)\n", name, best, cyc, clock_khz/1000);
  cudaFree(out);
}
int main(){
  int khz=1965000; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  run<0>("packed 3xFFMA2", khz);
  run<1>("packed 3xFFMA2 + FMNMX3", khz);
  run<2>("packed 3xFFMA2 + FMNMX", khz);
  run<3>("packed 3xFFMA2 + IADD", khz);
  run<4>("packed 3xFFMA2 + LDS", khz);
  run<10>("packed 3xFFMA2 + 2xFMNMX3", khz);
  run<11>("packed 3xFFMA2 + FSETP + @p MOV", khz);
  run<5>("scalar 6xFFMA", khz);
  run<6>("scalar 6xFFMA + FMNMX3", khz);
  run<7>("scalar 6xFFMA + 2xFMNMX", khz);
  run<8>("scalar 6xFFMA + IADD", khz);
  run<9>("scalar 6xFFMA + 2xFMNMX3", khz);
  run<12>("FMNMX3 alone", khz);
  run<13>("exact packed 3FADD2+FMUL2+2FFMA2+FMNMX3", khz);
  run<14>("exact scalar 6FADD+2FMUL+4FFMA+FMNMX3", khz);
  return 0;
}
