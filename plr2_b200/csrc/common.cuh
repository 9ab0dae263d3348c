// Shared device/host helpers for the plr2_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/plr2_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "plr2_b200 kernels are written for sm_100a (B200) only"
#endif

namespace plr2 {

// thread-local last-error text, set by the C-ABI wrappers
void set_error(const char *fmt, ...);

// One-time host set-up (cudaFuncSetAttribute, occupancy and SM-count queries) is PER DEVICE: a process may drive
// several GPUs (the Python layer accepts any device=).  Caches are small arrays indexed by the current device;
// entries are written with the same value by whichever thread gets there first (forward thread / autograd worker),
// through std::atomic so the race is benign.
constexpr int PLR2_MAX_DEVICES = 64;
inline int current_device() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= PLR2_MAX_DEVICES) d = 0;
  return d;
}
struct PerDeviceInt {
  std::atomic<int> v[PLR2_MAX_DEVICES];
  PerDeviceInt() { for (auto &x : v) x.store(0, std::memory_order_relaxed); }
  int get(int dev) const { return v[dev].load(std::memory_order_acquire); }
  void set(int dev, int val) { v[dev].store(val, std::memory_order_release); }
};
// SM count of the current device (cached per device)
int device_sm_count();
// Operand format of the INFERENCE tensor-core path (activation planes, packed weight planes): IEEE fp16 hi/lo
// (2 x 11 = 22 mantissa bits; default) or bf16 hi/lo (16 bits; PLR2_INFER_FMT=bf16, the round-1 format).  Same
// layout and MMA rate; the refiner's poses come out 10x closer to FP32, PoseNet's per-point outputs 3-4x
// (profiles/r2_precision_sweep*.txt).
// Values: activations are post-ReLU O(1e-3..1e2), weights O(1e-2): inside fp16's range; lo parts fall into fp16
// subnormals (absolute spacing 6e-8), which tcgen05 multiplies exactly.  The training path keeps bf16 planes
// (gradients span more than fp16's range).
bool infer_fp16();
// Products per MAC of conv6 on the inference path: 2 = X_hi (W_hi + W_lo) (default), 1 = X_hi W_hi (PLR2_CONV6_NPROD=1,
// experiment only: measured -0.40 ms of a 8.27 ms step - the layer turns shared-memory/L2 bound at 0.79 of the MMA
// peak instead of halving - while the refiner's pose error grows 20x to 1e-4 of its largest component: the weight
// rounding is the same for every point and does not average out in the pool)
int infer_conv6_nprod();

// number of kernels this library has launched (plr2_launch_count)
extern long long g_launch_count;

inline int check_launch(const char *what, int launches = 1) {
  g_launch_count += launches;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

// ---- packed FP32x2 arithmetic (Blackwell FADD2 / FMUL2 / FFMA2) -----------------------------
// Each op is IEEE round-to-nearest per half, i.e. bit-identical to the scalar add/mul/fma.
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
// 3-input min (FMNMX3): NaN operands are ignored like fminf
__device__ __forceinline__ float min3(float a, float b, float c) {
  float r;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// quaternion (w,x,y,z), already normalised -> rotation matrix, row-major, using exactly the
// expression order of lib/loss.py:17-25
__device__ __forceinline__ void quat_to_rot(float w, float x, float y, float z, float R[9]) {
  R[0] = 1.0f - 2.0f * (y * y + z * z);
  R[1] = 2.0f * x * y - 2.0f * w * z;
  R[2] = 2.0f * w * y + 2.0f * x * z;
  R[3] = 2.0f * x * y + 2.0f * z * w;
  R[4] = 1.0f - 2.0f * (x * x + z * z);
  R[5] = -2.0f * w * x + 2.0f * y * z;
  R[6] = -2.0f * w * y + 2.0f * x * z;
  R[7] = 2.0f * w * x + 2.0f * y * z;
  R[8] = 1.0f - 2.0f * (x * x + y * y);
}

}  // namespace plr2
