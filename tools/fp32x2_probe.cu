// Micro-probe: sustained throughput of packed FP32x2 vs scalar FP32 ops on B200 (sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32x2_probe tools/fp32x2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){u64 r; asm("mov.b64 %0,{%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ u64 fma2(u64 a,u64 b,u64 c){u64 r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;":"=l"(r):"l"(a),"l"(b),"l"(c)); return r;}
__device__ __forceinline__ u64 add2(u64 a,u64 b){u64 r; asm volatile("add.rn.f32x2 %0,%1,%2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ u64 mul2(u64 a,u64 b){u64 r; asm volatile("mul.rn.f32x2 %0,%1,%2;":"=l"(r):"l"(a),"l"(b)); return r;}
constexpr int NACC = 8, ITERS = 2048;
template <int MODE> __global__ void probe(float* out, float s) {
  u64 acc[NACC]; float sacc[2*NACC];
  for (int i=0;i<NACC;i++){acc[i]=pk(threadIdx.x*1e-3f+i, i*0.5f); sacc[2*i]=i+threadIdx.x; sacc[2*i+1]=i*2.f;}
  u64 b = pk(s, s*1.0001f), c = pk(s*0.5f, s*0.25f);
  float fb = s, fc = s*0.5f;
  for (int it=0; it<ITERS; ++it) {
#pragma unroll
    for (int i=0;i<NACC;i++) {
      if (MODE==0) acc[i]=fma2(acc[i], b, c);             // 3 distinct pairs
      if (MODE==1) acc[i]=fma2(b, b, acc[i]);             // a,a,c
      if (MODE==2) acc[i]=add2(acc[i], b);                // FADD2
      if (MODE==3) acc[i]=mul2(acc[i], acc[i]);           // FMUL2 a*a
      if (MODE==4) { asm volatile("fma.rn.f32 %0,%0,%1,%2;":"+f"(sacc[2*i]):"f"(fb),"f"(fc)); asm volatile("fma.rn.f32 %0,%0,%1,%2;":"+f"(sacc[2*i+1]):"f"(fb),"f"(fc)); }
      if (MODE==6) acc[i]=add2(acc[i], pk(fb, fb));      // FADD2 with scalar-broadcast operand (.F32 form)
      if (MODE==7) { u64 d=add2(b, pk(sacc[i], sacc[i])); acc[i]=fma2(d,d,acc[i]); }  // KNN-like: bcast add then fma
      if (MODE==8) { u64 d=add2(b, acc[(i+1)%NACC]); acc[i]=fma2(d,d,acc[i]); }         // same with a true pair operand
      if (MODE>=9 && MODE<=14) {
        acc[i]=fma2(b,b,acc[i]); acc[i]=fma2(c,c,acc[i]); acc[i]=fma2(b,c,acc[i]);
        float lo,hi; asm("mov.b64 {%0,%1},%2;":"=f"(lo),"=f"(hi):"l"(acc[i]));
        if (MODE==10) asm volatile("min.f32 %0,%0,%1,%2;":"+f"(sacc[i]):"f"(lo),"f"(hi));
        if (MODE==12) { int m=__float_as_int(sacc[i]); asm volatile("min.s32 %0,%0,%1;":"+r"(m):"r"(__float_as_int(lo))); asm volatile("min.s32 %0,%0,%1;":"+r"(m):"r"(__float_as_int(hi))); sacc[i]=__int_as_float(m); }
        if (MODE==13) { sacc[i]=__int_as_float(__vimin3_s32(__float_as_int(sacc[i]),__float_as_int(lo),__float_as_int(hi))); }
        if (MODE==14) { unsigned m=__float_as_uint(sacc[i]); asm volatile("min.u32 %0,%0,%1;":"+r"(m):"r"(__float_as_uint(lo))); asm volatile("min.u32 %0,%0,%1;":"+r"(m):"r"(__float_as_uint(hi))); sacc[i]=__uint_as_float(m); }
        if (MODE==11) { asm volatile("min.f32 %0,%0,%1;":"+f"(sacc[i]):"f"(lo)); asm volatile("min.f32 %0,%0,%1;":"+f"(sacc[i]):"f"(hi)); }
      }
      if (MODE==5) { acc[i]=add2(acc[i], b); acc[i]=fma2(acc[i],acc[i],c);}   // mixed like the KNN loop
    }
  }
  float r=0; for(int i=0;i<NACC;i++){float x,y; asm("mov.b64 {%0,%1},%2;":"=f"(x),"=f"(y):"l"(acc[i])); r+=x+y+sacc[2*i]+sacc[2*i+1];}
  out[blockIdx.x*blockDim.x+threadIdx.x]=r;
}
template <int MODE> void run(const char* name, int lanes_per_instr) {
  float* out; cudaMalloc(&out, 148*8*256*sizeof(float));
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  probe<MODE><<<148*8,256>>>(out,1.0001f); cudaDeviceSynchronize();
  cudaEventRecord(a); probe<MODE><<<148*8,256>>>(out,1.0001f); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms,a,b);
  double instr = (double)148*8*256*ITERS*NACC*(MODE==4||MODE==5||MODE==7||MODE==8?2:(MODE>=9?3:1));
  double lane_ops = instr*lanes_per_instr;    // fp32 lane-operations (FMA counts once)
  printf("%-28s %.3f ms  %.2f Tlane-op/s  (%.1f%% of 148*128*1.965GHz)\n", name, ms, lane_ops/ms/1e9, 100*lane_ops/ms/1e9/(148*128*1.965e-3));
  cudaFree(out);
}
int main(){
  run<4>("scalar FFMA", 1);
  run<0>("FFMA2 a,b,c", 2);
  run<1>("FFMA2 b,b,acc", 2);
  run<2>("FADD2 acc,b", 2);
  run<3>("FMUL2 a,a", 2);
  run<5>("FADD2+FFMA2 chain", 2);
  run<6>("FADD2 acc,bcast(.F32)", 2);
  run<7>("FADD2 bcast + FFMA2", 2);
  run<8>("FADD2 pair + FFMA2", 2);
  run<9>("3xFFMA2", 2);
  run<10>("3xFFMA2 + FMNMX3", 2);
  run<11>("3xFFMA2 + 2xFMNMX", 2);
  run<12>("3xFFMA2 + 2x min.s32", 2);
  run<13>("3xFFMA2 + vimin3_s32", 2);
  run<14>("3xFFMA2 + 2x min.u32", 2);
  return 0;
}
