// Kernels (3) and (5): quaternion -> rotation, mesh-point transform, ADD / ADD-S distance,
// confidence-weighted loss and its local Jacobians, best-vote selection and re-centring — fused so
// that neither the (N, M, 3) prediction tensor nor the KNN index tensor is ever materialised.
// Reference: lib/loss.py:12-69, lib/loss_refiner.py:12-62.
#include "knn_core.cuh"
#include <stdlib.h>
namespace plr2 {

// ---------------------------------------------------------------------------------------------
// Per-hypothesis pose: normalised quaternion, rotation (row-major) and translation.
struct Pose {
  float q[4];   // normalised (w,x,y,z)
  float inv_n;  // 1/||pred_r||
  float R[9];
  float s[3];   // translation added after rotation
};

__device__ __forceinline__ void make_pose(const float *__restrict__ r4, Pose &P) {
  const float a = r4[0], b = r4[1], c = r4[2], d = r4[3];
  const float n = sqrtf(a * a + b * b + c * c + d * d);  // torch.norm(pred_r, dim=2), lib/loss.py:15
  P.inv_n = 1.0f / n;
  P.q[0] = a / n; P.q[1] = b / n; P.q[2] = c / n; P.q[3] = d / n;
  quat_to_rot(P.q[0], P.q[1], P.q[2], P.q[3], P.R);
}

// Accumulated per-hypothesis sums over mesh points: distance, d/ds and d/dR.
struct Acc {
  float dis;
  float gs[3];
  float G[9];
};
__device__ __forceinline__ void acc_zero(Acc &A) {
  A.dis = 0.f;
#pragma unroll
  for (int i = 0; i < 3; ++i) A.gs[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 9; ++i) A.G[i] = 0.f;
}

// One (prediction, target) pair: adds ||p - t|| and, if JAC, the unit residual's outer products.
template <bool JAC>
__device__ __forceinline__ void acc_pair(Acc &A, float px, float py, float pz, float tx, float ty,
                                         float tz, float mx, float my, float mz) {
  const float dx = px - tx, dy = py - ty, dz = pz - tz;
  const float d2 = dx * dx + dy * dy + dz * dz;
  const float d = sqrtf(d2);
  A.dis += d;
  if (JAC) {
    const float inv = d > 0.f ? 1.0f / d : 0.f;  // torch: d||x||/dx = 0 at x = 0
    const float ux = dx * inv, uy = dy * inv, uz = dz * inv;
    A.gs[0] += ux; A.gs[1] += uy; A.gs[2] += uz;
    A.G[0] = fmaf(ux, mx, A.G[0]); A.G[1] = fmaf(ux, my, A.G[1]); A.G[2] = fmaf(ux, mz, A.G[2]);
    A.G[3] = fmaf(uy, mx, A.G[3]); A.G[4] = fmaf(uy, my, A.G[4]); A.G[5] = fmaf(uy, mz, A.G[5]);
    A.G[6] = fmaf(uz, mx, A.G[6]); A.G[7] = fmaf(uz, my, A.G[7]); A.G[8] = fmaf(uz, mz, A.G[8]);
  }
}

// Chain d(dis)/dR (G, already divided by M) through R(q_hat) and q_hat = r/||r|| to d/d pred_r.
// The partials are those of the exact expressions in quat_to_rot (lib/loss.py:17-25).
__device__ __forceinline__ void jac_quat(const Pose &P, const float G[9], float out[4]) {
  const float w = P.q[0], x = P.q[1], y = P.q[2], z = P.q[3];
  float gw = 2.f * (-z * G[1] + y * G[2] + z * G[3] - x * G[5] - y * G[6] + x * G[7]);
  float gx = 2.f * (y * G[1] + z * G[2] + y * G[3] - 2.f * x * G[4] - w * G[5] + z * G[6] + w * G[7] - 2.f * x * G[8]);
  float gy = 2.f * (-2.f * y * G[0] + x * G[1] + w * G[2] + x * G[3] + z * G[5] - w * G[6] + z * G[7] - 2.f * y * G[8]);
  float gz = 2.f * (-2.f * z * G[0] - w * G[1] + x * G[2] + w * G[3] - 2.f * z * G[4] + y * G[5] + x * G[6] + y * G[7]);
  const float dot = w * gw + x * gx + y * gy + z * gz;
  out[0] = (gw - w * dot) * P.inv_n;
  out[1] = (gx - x * dot) * P.inv_n;
  out[2] = (gy - y * dot) * P.inv_n;
  out[3] = (gz - z * dot) * P.inv_n;
}

__device__ __forceinline__ bool frame_is_sym(const int64_t *obj, const uint8_t *sym_lut, int num_obj, int f) {
  if (sym_lut == nullptr || obj == nullptr) return false;
  const long long o = obj[f];
  return o >= 0 && o < num_obj && sym_lut[o] != 0;
}

// ---------------------------------------------------------------------------------------------
// ADD (index-matched) distance: block = 32 hypotheses x 8 mesh slices.  Mesh and target points are
// staged as float4 so the 8 slices of a warp read one conflict-free 128 B row, broadcast over the
// 4 hypotheses sharing a slice.
constexpr int ADD_SLICES = 8;
constexpr int ADD_NPB = 32;     // hypotheses per block
constexpr int ADD_JT = 512;     // mesh points per smem tile

template <bool JAC>
__global__ void __launch_bounds__(ADD_SLICES * ADD_NPB)
add_loss_kernel(const float *__restrict__ pred_r, const float *__restrict__ pred_t,
                const float *__restrict__ target, const float *__restrict__ model,
                const float *__restrict__ points, const int64_t *__restrict__ obj,
                const uint8_t *__restrict__ sym_lut, int num_obj, int N, int M, int skip_sym,
                float *__restrict__ dis, float *__restrict__ jac_r, float *__restrict__ jac_t) {
  const int f = blockIdx.y;
  if (skip_sym && frame_is_sym(obj, sym_lut, num_obj, f)) return;  // handled by adds_loss_kernel
  __shared__ float4 sm[ADD_JT];
  __shared__ float4 st[ADD_JT];
  const int tid = threadIdx.x;
  const int slice = tid & (ADD_SLICES - 1);
  const int n = blockIdx.x * ADD_NPB + (tid >> 3);
  const bool live = n < N;
  const size_t fn = (size_t)f * N + (live ? n : 0);

  Pose P;
  make_pose(pred_r + fn * 4, P);
  if (points != nullptr) {
    P.s[0] = points[fn * 3 + 0] + pred_t[fn * 3 + 0];   // points + pred_t, lib/loss.py:37
    P.s[1] = points[fn * 3 + 1] + pred_t[fn * 3 + 1];
    P.s[2] = points[fn * 3 + 2] + pred_t[fn * 3 + 2];
  } else {
    P.s[0] = pred_t[fn * 3 + 0]; P.s[1] = pred_t[fn * 3 + 1]; P.s[2] = pred_t[fn * 3 + 2];
  }
  Acc A;
  acc_zero(A);
  const float *mb = model + (size_t)f * M * 3;
  const float *tb = target + (size_t)f * M * 3;
  for (int j0 = 0; j0 < M; j0 += ADD_JT) {
    const int jt = min(ADD_JT, M - j0);
    __syncthreads();
    for (int j = tid; j < jt; j += blockDim.x) {
      const float *m = mb + (size_t)(j0 + j) * 3;
      const float *t = tb + (size_t)(j0 + j) * 3;
      sm[j] = make_float4(__ldg(m), __ldg(m + 1), __ldg(m + 2), 0.f);
      st[j] = make_float4(__ldg(t), __ldg(t + 1), __ldg(t + 2), 0.f);
    }
    __syncthreads();
    for (int j = slice; j < jt; j += ADD_SLICES) {
      const float4 m = sm[j], t = st[j];
      const float px = fmaf(P.R[0], m.x, fmaf(P.R[1], m.y, fmaf(P.R[2], m.z, P.s[0])));
      const float py = fmaf(P.R[3], m.x, fmaf(P.R[4], m.y, fmaf(P.R[5], m.z, P.s[1])));
      const float pz = fmaf(P.R[6], m.x, fmaf(P.R[7], m.y, fmaf(P.R[8], m.z, P.s[2])));
      acc_pair<JAC>(A, px, py, pz, t.x, t.y, t.z, m.x, m.y, m.z);
    }
  }
  // reduce over the 8 slices (lanes differing in bits 0..2)
#pragma unroll
  for (int o = 1; o < ADD_SLICES; o <<= 1) {
    A.dis += __shfl_xor_sync(0xffffffffu, A.dis, o);
    if (JAC) {
#pragma unroll
      for (int i = 0; i < 3; ++i) A.gs[i] += __shfl_xor_sync(0xffffffffu, A.gs[i], o);
#pragma unroll
      for (int i = 0; i < 9; ++i) A.G[i] += __shfl_xor_sync(0xffffffffu, A.G[i], o);
    }
  }
  if (live && slice == 0) {
    const float invM = 1.0f / (float)M;
    dis[fn] = A.dis * invM;
    if (JAC) {
#pragma unroll
      for (int i = 0; i < 9; ++i) A.G[i] *= invM;
      float jq[4];
      jac_quat(P, A.G, jq);
#pragma unroll
      for (int i = 0; i < 4; ++i) jac_r[fn * 4 + i] = jq[i];
#pragma unroll
      for (int i = 0; i < 3; ++i) jac_t[fn * 3 + i] = A.gs[i] * invM;
    }
  }
}

// Inference variant (no Jacobians): thread = 4 hypotheses x 1/8 of the mesh, so that one broadcast
// LDS.128 pair (mesh point, target point) feeds 4 pair evaluations; sqrt.approx (<= 1 ulp-level error,
// far inside the rtol 1e-3 budget) instead of the IEEE sequence.
constexpr int ADDI_H = 4, ADDI_THREADS = 256;

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));     // one MUFU.SQRT (no denormal fix-up path)
  return r;
}

template <int ADDI_S>   // mesh slices per hypothesis group: 8, 4 or 2 (chosen so that the grid is ~one wave)
__global__ void __launch_bounds__(ADDI_THREADS)
add_dis_kernel(const float *__restrict__ pred_r, const float *__restrict__ pred_t,
               const float *__restrict__ target, const float *__restrict__ model,
               const float *__restrict__ points, const int64_t *__restrict__ obj,
               const uint8_t *__restrict__ sym_lut, int num_obj, int N, int M, int skip_sym,
               float *__restrict__ dis) {
  const int f = blockIdx.y;
  if (skip_sym && frame_is_sym(obj, sym_lut, num_obj, f)) return;  // handled by adds_loss_kernel
  __shared__ float4 sm[ADD_JT];
  __shared__ float4 st[ADD_JT];
  const int tid = threadIdx.x;
  constexpr int ADDI_NPB = (ADDI_THREADS / ADDI_S) * ADDI_H;     // hypotheses per block
  const int slice = tid & (ADDI_S - 1);
  const int n0 = blockIdx.x * ADDI_NPB + (tid / ADDI_S) * ADDI_H;
  // Hypotheses are processed as PAIRS in packed FP32x2 registers (FFMA2/FADD2: half the issue slots):
  // R2[p][i] = (R_{2p}[i], R_{2p+1}[i]) etc.; the mesh / target coordinates enter as broadcast operands.
  constexpr int HP = ADDI_H / 2;
  f32x2 R2[HP][9], s2[HP][3], acc2[HP];
#pragma unroll
  for (int p = 0; p < HP; ++p) {
    Pose P[2];
    float sf[2][3];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const size_t fn = (size_t)f * N + min(n0 + 2 * p + e, N - 1);
      make_pose(pred_r + fn * 4, P[e]);
#pragma unroll
      for (int i = 0; i < 3; ++i) sf[e][i] = points[fn * 3 + i] + pred_t[fn * 3 + i];   // lib/loss.py:37
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) R2[p][i] = pack2(P[0].R[i], P[1].R[i]);
#pragma unroll
    for (int i = 0; i < 3; ++i) s2[p][i] = pack2(sf[0][i], sf[1][i]);
    acc2[p] = pack2(0.f, 0.f);
  }
  const float *mb = model + (size_t)f * M * 3;
  const float *tb = target + (size_t)f * M * 3;
  for (int j0 = 0; j0 < M; j0 += ADD_JT) {
    const int jt = min(ADD_JT, M - j0);
    __syncthreads();
    for (int j = tid; j < jt; j += ADDI_THREADS) {
      const float *m = mb + (size_t)(j0 + j) * 3;
      const float *t = tb + (size_t)(j0 + j) * 3;
      sm[j] = make_float4(__ldg(m), __ldg(m + 1), __ldg(m + 2), 0.f);
      st[j] = make_float4(-__ldg(t), -__ldg(t + 1), -__ldg(t + 2), 0.f);   // negated: residual = pred + (-t)
    }
    __syncthreads();
#pragma unroll 2
    for (int j = slice; j < jt; j += ADDI_S) {
      const float4 m = sm[j], t = st[j];
      const f32x2 mx = pack2(m.x, m.x), my = pack2(m.y, m.y), mz = pack2(m.z, m.z);
      const f32x2 tx = pack2(t.x, t.x), ty = pack2(t.y, t.y), tz = pack2(t.z, t.z);
#pragma unroll
      for (int p = 0; p < HP; ++p) {
        const f32x2 dx = add2(fma2(R2[p][0], mx, fma2(R2[p][1], my, fma2(R2[p][2], mz, s2[p][0]))), tx);
        const f32x2 dy = add2(fma2(R2[p][3], mx, fma2(R2[p][4], my, fma2(R2[p][5], mz, s2[p][1]))), ty);
        const f32x2 dz = add2(fma2(R2[p][6], mx, fma2(R2[p][7], my, fma2(R2[p][8], mz, s2[p][2]))), tz);
        const f32x2 d2 = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
        float a, b;
        unpack2(d2, a, b);
        acc2[p] = add2(acc2[p], pack2(sqrt_approx(a), sqrt_approx(b)));
      }
    }
  }
  float acc[ADDI_H];
#pragma unroll
  for (int p = 0; p < HP; ++p) unpack2(acc2[p], acc[2 * p], acc[2 * p + 1]);
#pragma unroll
  for (int h = 0; h < ADDI_H; ++h) {
#pragma unroll
    for (int o = 1; o < ADDI_S; o <<= 1) acc[h] += __shfl_xor_sync(0xffffffffu, acc[h], o);
  }
  if (slice == 0) {
    const float invM = 1.0f / (float)M;
#pragma unroll
    for (int h = 0; h < ADDI_H; ++h)
      if (n0 + h < N) dis[(size_t)f * N + n0 + h] = acc[h] * invM;
  }
}

// ---------------------------------------------------------------------------------------------
// Block-wide sum of an Acc; result valid in thread 0.  `red` needs (blockDim/32)*13 floats.
template <bool JAC>
__device__ __forceinline__ void acc_block_reduce(Acc &A, float *red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  A.dis = warp_sum(A.dis);
  if (JAC) {
#pragma unroll
    for (int i = 0; i < 3; ++i) A.gs[i] = warp_sum(A.gs[i]);
#pragma unroll
    for (int i = 0; i < 9; ++i) A.G[i] = warp_sum(A.G[i]);
  }
  if (lane == 0) {
    red[warp * 13] = A.dis;
    if (JAC) {
#pragma unroll
      for (int i = 0; i < 3; ++i) red[warp * 13 + 1 + i] = A.gs[i];
#pragma unroll
      for (int i = 0; i < 9; ++i) red[warp * 13 + 4 + i] = A.G[i];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < nwarp; ++w) {
      A.dis += red[w * 13];
      if (JAC) {
#pragma unroll
        for (int i = 0; i < 3; ++i) A.gs[i] += red[w * 13 + 1 + i];
#pragma unroll
        for (int i = 0; i < 9; ++i) A.G[i] += red[w * 13 + 4 + i];
      }
    }
  }
}

// Stage the M target points of one frame as the SoA reference tile of the 1-NN scan.
// [A coalesced float4 read of the AoS rows scattered into the SoA arrays, with |r-c|^2 in a second pass, was measured
// SLOWER than this scalar form (all-symmetric 512 x 2300^2: 546 -> 565 us): the three strided loads per point already
// overlap across the loop iterations, the scatter's index arithmetic and the extra barrier do not pay.]
template <bool FAST>
__device__ __forceinline__ void stage_targets(const float *__restrict__ tb, int M, int Mp, float *sx,
                                              float *sy, float *sz, float *sw, float cx, float cy,
                                              float cz) {
  // Every CTA of a symmetric frame stages all M targets, so for short CTAs this loop is on the critical path: the loads
  // of STAGE_B points per thread are issued before the first one is used (ncu at 64 frames: 22 % of the stall samples
  // were long-scoreboard waits on one point's three loads at a time).
  constexpr int STAGE_B = 8;
  for (int j0 = threadIdx.x; j0 < Mp; j0 += STAGE_B * blockDim.x) {
    float x[STAGE_B], y[STAGE_B], z[STAGE_B];
#pragma unroll
    for (int b = 0; b < STAGE_B; ++b) {
      const int j = j0 + b * blockDim.x;
      x[b] = y[b] = z[b] = 0.f;
      if (j < M) { x[b] = __ldg(tb + 3 * j); y[b] = __ldg(tb + 3 * j + 1); z[b] = __ldg(tb + 3 * j + 2); }
    }
#pragma unroll
    for (int b = 0; b < STAGE_B; ++b) {
      const int j = j0 + b * blockDim.x;
      if (j >= Mp) break;
      float xx = x[b], yy = y[b], zz = z[b];
      if (j < M) {
        if (FAST) {
          xx -= cx; yy -= cy; zz -= cz;
          sw[j] = fmaf(zz, zz, fmaf(yy, yy, xx * xx));
        }
      } else {
        xx = yy = zz = FAST ? 0.f : CUDART_INF_F;
        if (FAST) sw[j] = CUDART_INF_F;
      }
      sx[j] = xx; sy[j] = yy; sz[j] = zz;
    }
  }
}

// ADD-S accumulation over the mesh points j in [j_beg, j_end) of one pose: transform, 1-NN over the
// staged targets, residual to the nearest target.  The slice is walked in rounds of blockDim.x lanes x
// QPT mesh points, QPT from {8,4,2,1} by binary decomposition of the per-lane count (M = 2300 on 128
// lanes: 8 + 8 + 2), like kernel (4).  Only the negated queries survive the scan in registers; the mesh
// point is re-read and re-transformed (same FMAs, same bits) for the residual.
__device__ __forceinline__ void adds_transform(const Pose &P, const float *__restrict__ mb, int j, float &mx,
                                               float &my, float &mz, float &px, float &py, float &pz) {
  mx = __ldg(mb + 3 * j); my = __ldg(mb + 3 * j + 1); mz = __ldg(mb + 3 * j + 2);
  px = fmaf(P.R[0], mx, fmaf(P.R[1], my, fmaf(P.R[2], mz, P.s[0])));
  py = fmaf(P.R[3], mx, fmaf(P.R[4], my, fmaf(P.R[5], mz, P.s[1])));
  pz = fmaf(P.R[6], mx, fmaf(P.R[7], my, fmaf(P.R[8], mz, P.s[2])));
}

template <bool FAST, bool JAC, int QPT>
__device__ __forceinline__ void adds_round(const Pose &P, const float *__restrict__ mb,
                                           const float *__restrict__ tb, int M, int Mp, int j0, int j_end,
                                           const float *sx, const float *sy, const float *sz,
                                           const float *sw, float cx, float cy, float cz, Acc &A) {
  const int stride = blockDim.x;
  // warps whose mesh points all lie beyond the slice skip the round (warp-uniform)
  if (j0 + (int)(threadIdx.x & ~31u) >= j_end) return;
  float nqx[QPT], nqy[QPT], nqz[QPT], best[QPT];
  int bchunk[QPT];
#pragma unroll
  for (int i = 0; i < QPT; ++i) {
    const int j = j0 + i * stride + threadIdx.x;
    float mx, my, mz, px, py, pz;
    adds_transform(P, mb, min(j, j_end - 1), mx, my, mz, px, py, pz);
    if (FAST) {
      nqx[i] = -2.f * (px - cx); nqy[i] = -2.f * (py - cy); nqz[i] = -2.f * (pz - cz);
    } else {
      nqx[i] = -px; nqy[i] = -py; nqz[i] = -pz;
    }
    best[i] = CUDART_INF_F;
    bchunk[i] = 0;
  }
  nn_scan<FAST, QPT>(sx, sy, sz, sw, Mp, 0, nqx, nqy, nqz, best, bchunk);
#pragma unroll
  for (int i = 0; i < QPT; ++i) {
    const int j = j0 + i * stride + threadIdx.x;
    if (j < j_end) {
      const int win = nn_resolve_smem4<FAST>(sx, sy, sz, sw, M, bchunk[i], best[i], nqx[i], nqy[i], nqz[i]);
      float mx, my, mz, px, py, pz;
      adds_transform(P, mb, j, mx, my, mz, px, py, pz);
      const float tx = __ldg(tb + 3 * win), ty = __ldg(tb + 3 * win + 1), tz = __ldg(tb + 3 * win + 2);
      acc_pair<JAC>(A, px, py, pz, tx, ty, tz, mx, my, mz);
    }
  }
}

template <bool FAST, bool JAC>
__device__ __forceinline__ void adds_accumulate(const Pose &P, const float *__restrict__ mb,
                                                const float *__restrict__ tb, int M, int Mp, int j_beg, int j_end,
                                                const float *sx, const float *sy, const float *sz,
                                                const float *sw, float cx, float cy, float cz, Acc &A) {
  const int stride = blockDim.x;
  int j = j_beg;
  while (j < j_end) {
    const int rem = (j_end - j + stride - 1) / stride;   // mesh points left per lane
    if (rem >= 8) {
      adds_round<FAST, JAC, 8>(P, mb, tb, M, Mp, j, j_end, sx, sy, sz, sw, cx, cy, cz, A);
      j += 8 * stride;
    } else if (rem >= 4) {
      adds_round<FAST, JAC, 4>(P, mb, tb, M, Mp, j, j_end, sx, sy, sz, sw, cx, cy, cz, A);
      j += 4 * stride;
    } else if (rem >= 2) {
      adds_round<FAST, JAC, 2>(P, mb, tb, M, Mp, j, j_end, sx, sy, sz, sw, cx, cy, cz, A);
      j += 2 * stride;
    } else {
      adds_round<FAST, JAC, 1>(P, mb, tb, M, Mp, j, j_end, sx, sy, sz, sw, cx, cy, cz, A);
      j += stride;
    }
  }
}

// ADD-S for symmetric frames of Loss (lib/loss.py:39-46): one CTA per (hypothesis n, frame).
constexpr int ADDS_THREADS = 128;
template <bool FAST, bool JAC>
__global__ void __launch_bounds__(ADDS_THREADS, 4)
adds_loss_kernel(const float *__restrict__ pred_r, const float *__restrict__ pred_t,
                 const float *__restrict__ target, const float *__restrict__ model,
                 const float *__restrict__ points, const int64_t *__restrict__ obj,
                 const uint8_t *__restrict__ sym_lut, int num_obj, int N, int M,
                 float *__restrict__ dis, float *__restrict__ jac_r, float *__restrict__ jac_t) {
  const int f = blockIdx.y;
  if (!frame_is_sym(obj, sym_lut, num_obj, f)) return;
  extern __shared__ __align__(16) float dyn[];
  __shared__ float red[(ADDS_THREADS / 32) * 13];
  const int Mp = knn_pad(M);
  float *sx = dyn, *sy = dyn + Mp, *sz = dyn + 2 * Mp, *sw = dyn + 3 * Mp;
  const float *tb = target + (size_t)f * M * 3;
  const float *mb = model + (size_t)f * M * 3;
  float cx = 0.f, cy = 0.f, cz = 0.f;
  if (FAST) { cx = __ldg(tb); cy = __ldg(tb + 1); cz = __ldg(tb + 2); }
  stage_targets<FAST>(tb, M, Mp, sx, sy, sz, sw, cx, cy, cz);
  const size_t fn = (size_t)f * N + blockIdx.x;
  Pose P;
  make_pose(pred_r + fn * 4, P);
  P.s[0] = points[fn * 3 + 0] + pred_t[fn * 3 + 0];
  P.s[1] = points[fn * 3 + 1] + pred_t[fn * 3 + 1];
  P.s[2] = points[fn * 3 + 2] + pred_t[fn * 3 + 2];
  __syncthreads();
  Acc A;
  acc_zero(A);
  adds_accumulate<FAST, JAC>(P, mb, tb, M, Mp, 0, M, sx, sy, sz, sw, cx, cy, cz, A);
  acc_block_reduce<JAC>(A, red);
  if (threadIdx.x == 0) {
    const float invM = 1.0f / (float)M;
    dis[fn] = A.dis * invM;
    if (JAC) {
#pragma unroll
      for (int i = 0; i < 9; ++i) A.G[i] *= invM;
      float jq[4];
      jac_quat(P, A.G, jq);
#pragma unroll
      for (int i = 0; i < 4; ++i) jac_r[fn * 4 + i] = jq[i];
#pragma unroll
      for (int i = 0; i < 3; ++i) jac_t[fn * 3 + i] = A.gs[i] * invM;
    }
  }
}

// loss[f] = mean_n (dis_n c_n - w log c_n)   (lib/loss.py:49); one CTA per frame, fixed order.
__global__ void __launch_bounds__(256)
loss_reduce_kernel(const float *__restrict__ dis, const float *__restrict__ pred_c, int N, float w,
                   float *__restrict__ loss) {
  __shared__ float red[8];
  const int f = blockIdx.x;
  float s = 0.f;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    const float c = pred_c[(size_t)f * N + n];
    s += dis[(size_t)f * N + n] * c - w * logf(c);
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    loss[f] = t / (float)N;
  }
}

// Backward combine: upstream grads of loss[f] and of the selected dis[f][which_max[f]].
__global__ void __launch_bounds__(256)
pose_loss_bwd_kernel(const float *__restrict__ g_loss, const float *__restrict__ g_dis_sel,
                     const int32_t *__restrict__ which_max, const float *__restrict__ pred_c,
                     const float *__restrict__ dis, const float *__restrict__ jac_r,
                     const float *__restrict__ jac_t, int F, int N, float w,
                     float *__restrict__ g_r, float *__restrict__ g_t, float *__restrict__ g_c) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)F * N) return;
  const int f = (int)(i / N), n = (int)(i % N);
  const float gl = g_loss ? g_loss[f] : 0.f;
  const float c = pred_c[i];
  float coef = gl * c / (float)N;
  if (g_dis_sel != nullptr && which_max != nullptr && which_max[f] == n) coef += g_dis_sel[f];
#pragma unroll
  for (int k = 0; k < 4; ++k) g_r[i * 4 + k] = coef * jac_r[i * 4 + k];
#pragma unroll
  for (int k = 0; k < 3; ++k) g_t[i * 3 + k] = coef * jac_t[i * 3 + k];
  if (g_c != nullptr) g_c[i] = gl * (dis[i] - w / c) / (float)N;
}

// ---------------------------------------------------------------------------------------------
// (5) best-vote selection + re-centring, lib/loss.py:52-69.  One CTA per frame.
__device__ __forceinline__ void recentre_rows(const float *__restrict__ src, float *__restrict__ dst,
                                              int rows, const float *R, const float *t) {
  for (int i = threadIdx.x; i < rows; i += blockDim.x) {
    const float a = src[3 * i] - t[0], b = src[3 * i + 1] - t[1], c = src[3 * i + 2] - t[2];
    // (x - t) . R  (row vector times ori_base, lib/loss.py:62,66)
    dst[3 * i + 0] = a * R[0] + b * R[3] + c * R[6];
    dst[3 * i + 1] = a * R[1] + b * R[4] + c * R[7];
    dst[3 * i + 2] = a * R[2] + b * R[5] + c * R[8];
  }
}

__global__ void __launch_bounds__(256)
pose_select_kernel(const float *__restrict__ pred_r, const float *__restrict__ pred_t,
                   const float *__restrict__ pred_c, const float *__restrict__ dis,
                   const float *__restrict__ target, const float *__restrict__ points, int N, int M,
                   int32_t *__restrict__ which_max, float *__restrict__ dis_sel,
                   float *__restrict__ new_points, float *__restrict__ new_target,
                   float *__restrict__ pose_out) {
  __shared__ float sval[8];
  __shared__ int sidx[8];
  __shared__ float sR[9], sT[3];
  const int f = blockIdx.x;
  const float *c = pred_c + (size_t)f * N;
  // argmax, lowest index on ties (torch.max over dim 1, lib/loss.py:53)
  float bv = -CUDART_INF_F;
  int bi = 0x7fffffff;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    const float v = c[n];
    if (v > bv || bi == 0x7fffffff) { bv = v; bi = n; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sval[threadIdx.x >> 5] = bv; sidx[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) {
      const float ov = sval[w];
      const int oi = sidx[w];
      if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
    }
    if (bi == 0x7fffffff) bi = 0;
    const size_t fn = (size_t)f * N + bi;
    Pose P;
    make_pose(pred_r + fn * 4, P);
#pragma unroll
    for (int i = 0; i < 9; ++i) sR[i] = P.R[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) sT[i] = pred_t[fn * 3 + i] + points[fn * 3 + i];  // lib/loss.py:57
    if (which_max) which_max[f] = bi;
    if (dis_sel && dis) dis_sel[f] = dis[fn];
    if (pose_out) {
#pragma unroll
      for (int i = 0; i < 9; ++i) pose_out[(size_t)f * 12 + i] = P.R[i];
#pragma unroll
      for (int i = 0; i < 3; ++i) pose_out[(size_t)f * 12 + 9 + i] = sT[i];
    }
  }
  __syncthreads();
  if (new_points) recentre_rows(points + (size_t)f * N * 3, new_points + (size_t)f * N * 3, N, sR, sT);
  if (new_target) recentre_rows(target + (size_t)f * M * 3, new_target + (size_t)f * M * 3, M, sR, sT);
}

// ---------------------------------------------------------------------------------------------
// Refiner loss, lib/loss_refiner.py:12-62: one pose per frame, up to 32 CTAs per frame: CTA s owns mesh
// points [s*slice, (s+1)*slice) (each CTA stages all M targets for the ADD-S scan) and the matching
// slices of the re-centring; per-CTA partial sums meet in the caller's workspace and are combined in
// slice order by the last CTA to arrive (deterministic).  [A thread-block-cluster/DSMEM version was
// measured first: 64 clusters of 8 cost ~75 us per launch regardless of the work inside.]
constexpr int REFINE_RS = 3;                       // reference ranges scanned concurrently per query
constexpr int REFINE_QT = 96;                      // query threads per range (3 warps)
constexpr int REFINE_THREADS = REFINE_RS * REFINE_QT;
// Queries per thread: 1 when few frames are in flight (maximise warps: the whole launch is a handful of
// CTAs per SM), 4 for large batches - with one query per thread the scan issues one LDS.128 per 1.5 FFMA2
// and sits on the shared-memory pipe (ncu: L1 61 % busy, FMA 26 %), four queries share each load.
constexpr int REFINE_QPT_MAX = 4;
constexpr int REFINE_PASS1 = REFINE_QT;            // queries per CTA pass at one query per thread
constexpr int REFINE_MAX_SPLIT = 32;               // CTAs per frame

// ADD-S accumulation for the refiner: thread (r, qt) scans reference range r for REFINE_QPT queries;
// the per-range candidates meet in shared memory and range-0 threads finish (lowest range wins ties,
// so the lowest reference index still wins, knn.cu:158).
template <bool FAST, bool JAC, int REFINE_QPT>
__device__ __forceinline__ void adds_accumulate_split(const Pose &P, const float *__restrict__ mb,
                                                      const float *__restrict__ tb, int M, int Mp, int j_beg,
                                                      int j_end, const float *sx, const float *sy, const float *sz,
                                                      const float *sw, float cx, float cy, float cz,
                                                      float (*cand_d)[REFINE_QT * REFINE_QPT], int (*cand_c)[REFINE_QT * REFINE_QPT], Acc &A) {
  constexpr int REFINE_PASS = REFINE_QT * REFINE_QPT;
  const int lane = threadIdx.x & 31;
  const int r = threadIdx.x / REFINE_QT, qt = threadIdx.x - r * REFINE_QT;
  const int nchunk = Mp / KNN_CHUNK;
  const int c_beg = (int)((long long)nchunk * r / REFINE_RS), c_end = (int)((long long)nchunk * (r + 1) / REFINE_RS);
  for (int j0 = j_beg; j0 < j_end; j0 += REFINE_PASS) {
    float nqx[REFINE_QPT], nqy[REFINE_QPT], nqz[REFINE_QPT], best[REFINE_QPT];
    int bchunk[REFINE_QPT];
#pragma unroll
    for (int i = 0; i < REFINE_QPT; ++i) {
      const int j = j0 + i * REFINE_QT + qt;
      float mx, my, mz, px, py, pz;
      adds_transform(P, mb, min(j, j_end - 1), mx, my, mz, px, py, pz);
      if (FAST) {
        nqx[i] = -2.f * (px - cx); nqy[i] = -2.f * (py - cy); nqz[i] = -2.f * (pz - cz);
      } else {
        nqx[i] = -px; nqy[i] = -py; nqz[i] = -pz;
      }
      best[i] = CUDART_INF_F;
      bchunk[i] = c_beg;
    }
    if (j0 + (qt & ~31) < j_end)        // warp-uniform: warps whose queries all lie beyond the slice skip the scan
      nn_scan<FAST, REFINE_QPT>(sx + c_beg * KNN_CHUNK, sy + c_beg * KNN_CHUNK, sz + c_beg * KNN_CHUNK,
                                sw + (FAST ? c_beg * KNN_CHUNK : 0), (c_end - c_beg) * KNN_CHUNK, c_beg, nqx, nqy, nqz,
                                best, bchunk);
#pragma unroll
    for (int i = 0; i < REFINE_QPT; ++i) {
      cand_d[r][i * REFINE_QT + qt] = best[i];
      cand_c[r][i * REFINE_QT + qt] = bchunk[i];
    }
    __syncthreads();
    if (r == 0) {
#pragma unroll
      for (int i = 0; i < REFINE_QPT; ++i) {
        const int j = j0 + i * REFINE_QT + qt;
        if (j < j_end) {
          float b = best[i];
          int bc = bchunk[i];
#pragma unroll
          for (int rr = 1; rr < REFINE_RS; ++rr) {
            const float d = cand_d[rr][i * REFINE_QT + qt];
            if (d < b) { b = d; bc = cand_c[rr][i * REFINE_QT + qt]; }
          }
          const int win = nn_resolve_smem<FAST>(sx, sy, sz, sw, M, bc, b, nqx[i], nqy[i], nqz[i], lane);
          const float tx = __ldg(tb + 3 * win), ty = __ldg(tb + 3 * win + 1), tz = __ldg(tb + 3 * win + 2);
          float mx, my, mz, px, py, pz;      // re-read and re-transform (same FMAs, same bits) instead of keeping them live
          adds_transform(P, mb, j, mx, my, mz, px, py, pz);
          acc_pair<JAC>(A, px, py, pz, tx, ty, tz, mx, my, mz);
        }
      }
    }
    __syncthreads();
  }
}

template <bool FAST, bool JAC, int QPT>
__global__ void __launch_bounds__(REFINE_THREADS, (QPT > 1 || (FAST && JAC)) ? 2 : 3)
refine_loss_kernel(const float *__restrict__ pred_r, const float *__restrict__ pred_t,
                   const float *__restrict__ target, const float *__restrict__ model,
                   const float *__restrict__ points, const int64_t *__restrict__ obj,
                   const uint8_t *__restrict__ sym_lut, int num_obj, int N, int M, int slice,
                   unsigned int *__restrict__ counters, float *__restrict__ partials,
                   float *__restrict__ dis, float *__restrict__ new_points,
                   float *__restrict__ new_target, float *__restrict__ jac_r,
                   float *__restrict__ jac_t, float *__restrict__ pose_out) {
  extern __shared__ __align__(16) float dyn[];
  __shared__ float red[(REFINE_THREADS / 32) * 13];
  __shared__ int s_last;
  __shared__ float cand_d[REFINE_RS][REFINE_QT * QPT];
  __shared__ int cand_c[REFINE_RS][REFINE_QT * QPT];
  const int S = gridDim.x, s = blockIdx.x;
  const int f = blockIdx.y;
  const float *tb = target + (size_t)f * M * 3;
  const float *mb = model + (size_t)f * M * 3;
  Pose P;
  make_pose(pred_r + (size_t)f * 4, P);
  P.s[0] = pred_t[(size_t)f * 3]; P.s[1] = pred_t[(size_t)f * 3 + 1]; P.s[2] = pred_t[(size_t)f * 3 + 2];
  const int j_beg = min(M, s * slice), j_end = min(M, j_beg + slice);
  Acc A;
  acc_zero(A);
  if (frame_is_sym(obj, sym_lut, num_obj, f)) {
    const int Mp = knn_pad(M);
    float *sx = dyn, *sy = dyn + Mp, *sz = dyn + 2 * Mp, *sw = dyn + 3 * Mp;
    float cx = 0.f, cy = 0.f, cz = 0.f;
    if (FAST) { cx = __ldg(tb); cy = __ldg(tb + 1); cz = __ldg(tb + 2); }
    stage_targets<FAST>(tb, M, Mp, sx, sy, sz, sw, cx, cy, cz);
    __syncthreads();
    adds_accumulate_split<FAST, JAC, QPT>(P, mb, tb, M, Mp, j_beg, j_end, sx, sy, sz, sw, cx, cy, cz, cand_d, cand_c, A);
  } else {
    for (int j = j_beg + threadIdx.x; j < j_end; j += blockDim.x) {
      const float mx = __ldg(mb + 3 * j), my = __ldg(mb + 3 * j + 1), mz = __ldg(mb + 3 * j + 2);
      const float px = fmaf(P.R[0], mx, fmaf(P.R[1], my, fmaf(P.R[2], mz, P.s[0])));
      const float py = fmaf(P.R[3], mx, fmaf(P.R[4], my, fmaf(P.R[5], mz, P.s[1])));
      const float pz = fmaf(P.R[6], mx, fmaf(P.R[7], my, fmaf(P.R[8], mz, P.s[2])));
      acc_pair<JAC>(A, px, py, pz, __ldg(tb + 3 * j), __ldg(tb + 3 * j + 1), __ldg(tb + 3 * j + 2), mx, my, mz);
    }
  }
  acc_block_reduce<JAC>(A, red);
  // Per-CTA partial sums go to the workspace; the CTA that arrives LAST for this frame adds them up
  // in slice order (deterministic) and writes the frame's outputs, then re-arms the counter.
  if (threadIdx.x == 0) {
    float *pp = partials + ((size_t)f * S + s) * 13;
    pp[0] = A.dis;
#pragma unroll
    for (int i = 0; i < 3; ++i) pp[1 + i] = JAC ? A.gs[i] : 0.f;
#pragma unroll
    for (int i = 0; i < 9; ++i) pp[4 + i] = JAC ? A.G[i] : 0.f;
    __threadfence();
    s_last = atomicAdd(counters + f, 1u) == (unsigned)(S - 1);
  }
  __syncthreads();
  if (s_last && threadIdx.x < 13) {
    // lane c sums component c of the S partial records in slice order (deterministic); the loads of a lane are
    // independent, so they are issued eight at a time instead of one L2 round trip per slice
    __threadfence();
    const float *rp = partials + (size_t)f * S * 13 + threadIdx.x;
    float tot = 0.f;
    for (int r0 = 0; r0 < S; r0 += 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = r0 + u < S ? __ldcg(rp + (size_t)(r0 + u) * 13) : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u) tot += v[u];
    }
    red[threadIdx.x] = tot;
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    A.dis = red[0];
#pragma unroll
    for (int i = 0; i < 3; ++i) A.gs[i] = red[1 + i];
#pragma unroll
    for (int i = 0; i < 9; ++i) A.G[i] = red[4 + i];
    counters[f] = 0;
    const float invM = 1.0f / (float)M;
    dis[f] = A.dis * invM;
    if (JAC) {
#pragma unroll
      for (int i = 0; i < 9; ++i) A.G[i] *= invM;
      float jq[4];
      jac_quat(P, A.G, jq);
#pragma unroll
      for (int i = 0; i < 4; ++i) jac_r[(size_t)f * 4 + i] = jq[i];
#pragma unroll
      for (int i = 0; i < 3; ++i) jac_t[(size_t)f * 3 + i] = A.gs[i] * invM;
    }
    if (pose_out) {
#pragma unroll
      for (int i = 0; i < 9; ++i) pose_out[(size_t)f * 12 + i] = P.R[i];
#pragma unroll
      for (int i = 0; i < 3; ++i) pose_out[(size_t)f * 12 + 9 + i] = P.s[i];
    }
  }
  // re-centring (lib/loss_refiner.py:50-59): every CTA takes its share of the rows
  if (new_points) {
    const int per = (N + S - 1) / S, a = min(N, s * per), b = min(N, a + per);
    recentre_rows(points + ((size_t)f * N + a) * 3, new_points + ((size_t)f * N + a) * 3, b - a, P.R, P.s);
  }
  if (new_target) {
    const int per = (M + S - 1) / S, a = min(M, s * per), b = min(M, a + per);
    recentre_rows(tb + (size_t)a * 3, new_target + ((size_t)f * M + a) * 3, b - a, P.R, P.s);
  }
}

// Pose composition across refinement stages (kernel (5), tools/train.py:141-145 hand-off): the stage
// pose (Rd, td) lives in the frame re-centred by the accumulated pose (R, t), x_rc = R^T (x - t), hence
//   R <- R Rd,   t <- R td + t.
__global__ void __launch_bounds__(128)
pose_compose_kernel(float *__restrict__ acc, const float *__restrict__ delta, int frames) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= frames) return;
  float R[9], t[3], D[9], d[3];
#pragma unroll
  for (int i = 0; i < 9; ++i) { R[i] = acc[(size_t)f * 12 + i]; D[i] = delta[(size_t)f * 12 + i]; }
#pragma unroll
  for (int i = 0; i < 3; ++i) { t[i] = acc[(size_t)f * 12 + 9 + i]; d[i] = delta[(size_t)f * 12 + 9 + i]; }
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      acc[(size_t)f * 12 + 3 * r + c] = R[3 * r] * D[c] + R[3 * r + 1] * D[3 + c] + R[3 * r + 2] * D[6 + c];
    acc[(size_t)f * 12 + 9 + r] = R[3 * r] * d[0] + R[3 * r + 1] * d[1] + R[3 * r + 2] * d[2] + t[r];
  }
}

static int set_dyn_smem(const void *fn, size_t bytes) {
  if (bytes <= 48 * 1024) return 0;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(smem=%zu): %s", bytes, cudaGetErrorString(e)); return (int)e; }
  return 0;
}

}  // namespace plr2

using namespace plr2;

#define PLR2_LAUNCH_ADDS(KERNEL, FASTV, JACV, GRID, BLOCK, SMEM, ST, ...)                      \
  do {                                                                                           \
    auto kfn = KERNEL<FASTV, JACV>;                                                              \
    int rc_ = set_dyn_smem((const void *)kfn, SMEM);                                             \
    if (rc_) return rc_;                                                                         \
    kfn<<<GRID, BLOCK, SMEM, ST>>>(__VA_ARGS__);                                                 \
  } while (0)

#define PLR2_LAUNCH_ADDS3(KERNEL, FASTV, JACV, QV, GRID, BLOCK, SMEM, ST, ...)                 \
  do {                                                                                           \
    auto kfn = KERNEL<FASTV, JACV, QV>;                                                          \
    int rc_ = set_dyn_smem((const void *)kfn, SMEM);                                             \
    if (rc_) return rc_;                                                                         \
    kfn<<<GRID, BLOCK, SMEM, ST>>>(__VA_ARGS__);                                                 \
  } while (0)

extern "C" int plr2_pose_loss_fwd(const float *pred_r, const float *pred_t, const float *pred_c,
                                  const float *target, const float *model_points, const float *points,
                                  const int64_t *obj, const uint8_t *sym_lut, int num_obj, int frames,
                                  int n_points, int n_mesh, float w, int refine, int knn_mode,
                                  float *dis, float *loss, float *jac_r, float *jac_t, void *stream) {
  if (frames < 0 || n_points <= 0 || n_mesh <= 0 || frames > 65535) {
    set_error("plr2_pose_loss_fwd: bad shape F=%d N=%d M=%d", frames, n_points, n_mesh);
    return PLR2_E_SHAPE;
  }
  if (frames == 0) return 0;
  if (!pred_r || !pred_t || !target || !model_points || !points || !dis || (loss && !pred_c) ||
      ((jac_r == nullptr) != (jac_t == nullptr))) {
    set_error("plr2_pose_loss_fwd: null pointer");
    return PLR2_E_NULLPTR;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const bool jac = jac_r != nullptr;
  const bool may_sym = !refine && sym_lut != nullptr && obj != nullptr;
  dim3 g1((n_points + ADD_NPB - 1) / ADD_NPB, frames);
  if (jac)
    add_loss_kernel<true><<<g1, ADD_SLICES * ADD_NPB, 0, st>>>(pred_r, pred_t, target, model_points, points, obj, sym_lut, num_obj, n_points, n_mesh, may_sym ? 1 : 0, dis, jac_r, jac_t);
  else {
    // Mesh slices per hypothesis group, S in {16, 8, 4, 2}: a block covers (256 / S) * 4 hypotheses, ~3 blocks are
    // resident per SM.  Cost model fitted to measurements (64 frames: 121 / 131 / 131 / 143 us for S = 16 / 8 / 4 / 2;
    // 512 frames: 795 / 755 / 737 / 738): waves(S) / S for the wave quantisation, times (1 + 0.02 S) for the mesh tile
    // every block re-stages; grids of many waves take S = 4.
    const int slots = 3 * device_sm_count();
    auto blocks = [&](int S) { const int npb = (ADDI_THREADS / S) * ADDI_H; return (long long)((n_points + npb - 1) / npb) * frames; };
#define ADDI_LAUNCH(S) do { const int npb = (ADDI_THREADS / S) * ADDI_H; dim3 gi((n_points + npb - 1) / npb, frames);   \
      add_dis_kernel<S><<<gi, ADDI_THREADS, 0, st>>>(pred_r, pred_t, target, model_points, points, obj, sym_lut, num_obj,  \
                                                     n_points, n_mesh, may_sym ? 1 : 0, dis); } while (0)
    static int force_s = -1;           // PLR2_ADDI_S: tuning experiments only
    if (force_s < 0) { const char *e = getenv("PLR2_ADDI_S"); force_s = e ? atoi(e) : 0; }
    if (force_s == 16) ADDI_LAUNCH(16);
    else if (force_s == 8) ADDI_LAUNCH(8);
    else if (force_s == 4) ADDI_LAUNCH(4);
    else if (force_s == 2) ADDI_LAUNCH(2);
    else {
      int best_s = 4;
      if (blocks(4) < 4LL * slots) {
        double best_c = 1e30;
        for (int S = 16; S >= 2; S >>= 1) {
          const double c = (double)((blocks(S) + slots - 1) / slots) / S * (1.0 + 0.02 * S);
          if (c < best_c) { best_c = c; best_s = S; }
        }
      }
      if (best_s == 16) ADDI_LAUNCH(16);
      else if (best_s == 8) ADDI_LAUNCH(8);
      else if (best_s == 4) ADDI_LAUNCH(4);
      else ADDI_LAUNCH(2);
    }
#undef ADDI_LAUNCH
  }
  if (may_sym) {
    if (n_mesh > 8192) { set_error("plr2_pose_loss_fwd: ADD-S needs n_mesh <= 8192 (got %d)", n_mesh); return PLR2_E_UNSUPPORTED; }
    const bool fast = knn_mode == PLR2_KNN_FAST;
    const size_t smem = (size_t)knn_pad(n_mesh) * sizeof(float) * (fast ? 4 : 3);
    dim3 g2(n_points, frames);
#define ARGS pred_r, pred_t, target, model_points, points, obj, sym_lut, num_obj, n_points, n_mesh, dis, jac_r, jac_t
    if (fast && jac) PLR2_LAUNCH_ADDS(adds_loss_kernel, true, true, g2, ADDS_THREADS, smem, st, ARGS);
    else if (fast) PLR2_LAUNCH_ADDS(adds_loss_kernel, true, false, g2, ADDS_THREADS, smem, st, ARGS);
    else if (jac) PLR2_LAUNCH_ADDS(adds_loss_kernel, false, true, g2, ADDS_THREADS, smem, st, ARGS);
    else PLR2_LAUNCH_ADDS(adds_loss_kernel, false, false, g2, ADDS_THREADS, smem, st, ARGS);
#undef ARGS
  }
  if (loss) loss_reduce_kernel<<<frames, 256, 0, st>>>(dis, pred_c, n_points, w, loss);
  return check_launch("plr2_pose_loss_fwd", 1 + (may_sym ? 1 : 0) + (loss ? 1 : 0));
}

extern "C" int plr2_pose_loss_bwd(const float *g_loss, const float *g_dis_sel, const int32_t *which_max,
                                  const float *pred_c, const float *dis, const float *jac_r,
                                  const float *jac_t, int frames, int n_points, float w,
                                  float *g_pred_r, float *g_pred_t, float *g_pred_c, void *stream) {
  if (frames < 0 || n_points <= 0) { set_error("plr2_pose_loss_bwd: bad shape"); return PLR2_E_SHAPE; }
  if (frames == 0) return 0;
  if (!pred_c || !dis || !jac_r || !jac_t || !g_pred_r || !g_pred_t) { set_error("plr2_pose_loss_bwd: null pointer"); return PLR2_E_NULLPTR; }
  const size_t tot = (size_t)frames * n_points;
  pose_loss_bwd_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      g_loss, g_dis_sel, which_max, pred_c, dis, jac_r, jac_t, frames, n_points, w, g_pred_r, g_pred_t, g_pred_c);
  return check_launch("plr2_pose_loss_bwd");
}

extern "C" int plr2_pose_select(const float *pred_r, const float *pred_t, const float *pred_c,
                                const float *dis, const float *target, const float *points, int frames,
                                int n_points, int n_mesh, int32_t *which_max, float *dis_sel,
                                float *new_points, float *new_target, float *pose_out, void *stream) {
  if (frames < 0 || n_points <= 0 || n_mesh < 0) { set_error("plr2_pose_select: bad shape"); return PLR2_E_SHAPE; }
  if (frames == 0) return 0;
  if (!pred_r || !pred_t || !pred_c || !points || (new_target && !target)) { set_error("plr2_pose_select: null pointer"); return PLR2_E_NULLPTR; }
  pose_select_kernel<<<frames, 256, 0, (cudaStream_t)stream>>>(pred_r, pred_t, pred_c, dis, target, points, n_points, n_mesh, which_max, dis_sel, new_points, new_target, pose_out);
  return check_launch("plr2_pose_select");
}

static int refine_split(int n_mesh, int qpt = 1) {
  const int pass = REFINE_PASS1 * qpt;
  int S = (n_mesh + pass - 1) / pass;
  return S > REFINE_MAX_SPLIT ? REFINE_MAX_SPLIT : (S < 1 ? 1 : S);
}

// workspace: [frames] arrival counters (u32, padded to 128 B) then [frames][S][13] partial sums
extern "C" size_t plr2_refine_loss_workspace_bytes(int frames, int n_mesh) {
  if (frames <= 0 || n_mesh <= 0) return 0;
  const size_t cnt = ((size_t)frames * sizeof(unsigned int) + 127) / 128 * 128;
  return cnt + (size_t)frames * refine_split(n_mesh) * 13 * sizeof(float);
}

extern "C" int plr2_refine_loss_fwd(const float *pred_r, const float *pred_t, const float *target,
                                    const float *model_points, const float *points, const int64_t *obj,
                                    const uint8_t *sym_lut, int num_obj, int frames, int n_points,
                                    int n_mesh, int knn_mode, float *dis, float *new_points,
                                    float *new_target, float *jac_r, float *jac_t, float *pose_out,
                                    void *workspace, size_t workspace_bytes, void *stream) {
  if (frames < 0 || n_points < 0 || n_mesh <= 0 || frames > 65535) { set_error("plr2_refine_loss_fwd: bad shape"); return PLR2_E_SHAPE; }
  if (frames == 0) return 0;
  if (!pred_r || !pred_t || !target || !model_points || !dis || (new_points && !points) ||
      ((jac_r == nullptr) != (jac_t == nullptr))) { set_error("plr2_refine_loss_fwd: null pointer"); return PLR2_E_NULLPTR; }
  if (n_mesh > 8192) { set_error("plr2_refine_loss_fwd: n_mesh <= 8192 required (got %d)", n_mesh); return PLR2_E_UNSUPPORTED; }
  if (!workspace || workspace_bytes < plr2_refine_loss_workspace_bytes(frames, n_mesh) ||
      (reinterpret_cast<uintptr_t>(workspace) & 15)) {
    set_error("plr2_refine_loss_fwd: workspace missing, misaligned or smaller than plr2_refine_loss_workspace_bytes");
    return PLR2_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const bool fast = knn_mode == PLR2_KNN_FAST, jac = jac_r != nullptr;
  const bool may_sym = sym_lut != nullptr && obj != nullptr;
  const size_t smem = may_sym ? (size_t)knn_pad(n_mesh) * sizeof(float) * (fast ? 4 : 3) : 0;
  // queries per thread: 4 once one-query CTAs alone would fill the GPU eight times over, 2 for medium batches (measured at
  // 64 frames, 5 of 16 symmetric: 61 / 51 / 65 us for 1 / 2 / 4), 1 for a handful of frames (most CTAs, shortest critical path)
  int qpt = (long long)frames * refine_split(n_mesh, 1) >= 8LL * 3 * device_sm_count() ? REFINE_QPT_MAX : (frames >= 24 ? 2 : 1);
  static int force_q = -1;             // PLR2_REFINE_QPT: tuning experiments only
  if (force_q < 0) { const char *e = getenv("PLR2_REFINE_QPT"); force_q = e ? atoi(e) : 0; }
  if (force_q == 1 || force_q == 2 || force_q == 4) qpt = force_q;
  const int S = refine_split(n_mesh, qpt);
  int slice = (n_mesh + S - 1) / S;
  slice = (slice + 31) / 32 * 32;               // warp granularity
  const size_t cnt_bytes = ((size_t)frames * sizeof(unsigned int) + 127) / 128 * 128;
  unsigned int *counters = reinterpret_cast<unsigned int *>(workspace);
  float *partials = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(workspace) + cnt_bytes);
  cudaError_t e = cudaMemsetAsync(counters, 0, cnt_bytes, st);
  if (e != cudaSuccess) { set_error("plr2_refine_loss_fwd: memset: %s", cudaGetErrorString(e)); return (int)e; }
  dim3 grid(S, frames);
#define ARGS pred_r, pred_t, target, model_points, points, obj, sym_lut, num_obj, n_points, n_mesh, slice, counters, partials, dis, new_points, new_target, jac_r, jac_t, pose_out
#define REFINE_LAUNCH(FASTV, JACV)                                                               \
  do {                                                                                           \
    if (qpt == 1) PLR2_LAUNCH_ADDS3(refine_loss_kernel, FASTV, JACV, 1, grid, REFINE_THREADS, smem, st, ARGS);                \
    else if (qpt == 2) PLR2_LAUNCH_ADDS3(refine_loss_kernel, FASTV, JACV, 2, grid, REFINE_THREADS, smem, st, ARGS);           \
    else PLR2_LAUNCH_ADDS3(refine_loss_kernel, FASTV, JACV, REFINE_QPT_MAX, grid, REFINE_THREADS, smem, st, ARGS);            \
  } while (0)
  if (fast && jac) REFINE_LAUNCH(true, true);
  else if (fast) REFINE_LAUNCH(true, false);
  else if (jac) REFINE_LAUNCH(false, true);
  else REFINE_LAUNCH(false, false);
#undef REFINE_LAUNCH
#undef ARGS
  return check_launch("plr2_refine_loss_fwd");
}

extern "C" int plr2_pose_compose(float *pose_acc, const float *pose_delta, int frames, void *stream) {
  if (frames < 0) { set_error("plr2_pose_compose: bad shape"); return PLR2_E_SHAPE; }
  if (frames == 0) return 0;
  if (!pose_acc || !pose_delta) { set_error("plr2_pose_compose: null pointer"); return PLR2_E_NULLPTR; }
  pose_compose_kernel<<<(frames + 127) / 128, 128, 0, (cudaStream_t)stream>>>(pose_acc, pose_delta, frames);
  return check_launch("plr2_pose_compose");
}

// ---------------------------------------------------------------------------------------------
// Refiner loss without caller-provided scratch (no counterpart in the reference; added
// for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  The per-frame arrival counters and per-CTA partial sums of refine_loss_kernel live in
// a file-scope device array of fixed size; every GPU that loads the library holds these 417 KB until exit.
// In HOST code the array's name is only a stand-in: its device address is obtained once per device, with that
// device current, by cudaGetSymbolAddress, and that pointer is what the kernel receives.
constexpr int STATIC_REFINE_MAX_FRAMES = PLR2_REFINE_LOSS_STATIC_MAX_FRAMES;
constexpr int STATIC_REFINE_MAX_MESH = 8192;                               // limit of refine_loss_kernel itself
constexpr size_t STATIC_REFINE_BYTES =
    ((size_t)STATIC_REFINE_MAX_FRAMES * sizeof(unsigned int) + 127) / 128 * 128 +
    (size_t)STATIC_REFINE_MAX_FRAMES * REFINE_MAX_SPLIT * 13 * sizeof(float);
static __device__ __align__(256) unsigned char g_refine_static_scratch[STATIC_REFINE_BYTES];

// device address of the array on the current device (cached per device; racing initialisers store the same value)
static int refine_static_scratch_ptr(void **out) {
  static std::atomic<void *> cache[plr2::PLR2_MAX_DEVICES];
  const int dev = plr2::current_device();
  void *p = cache[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    cudaError_t e = cudaGetSymbolAddress(&p, g_refine_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      plr2::set_error("plr2_refine_loss_fwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : PLR2_E_WORKSPACE;
    }
    cache[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}

extern "C" size_t plr2_refine_loss_static_scratch_bytes(void) { return STATIC_REFINE_BYTES; }

extern "C" int plr2_refine_loss_fwd_static(const float *pred_r, const float *pred_t, const float *target,
                                           const float *model_points, const float *points, const int64_t *obj,
                                           const uint8_t *sym_lut, int num_obj, int frames, int n_points,
                                           int n_mesh, int knn_mode, float *dis, float *new_points,
                                           float *new_target, float *jac_r, float *jac_t, float *pose_out,
                                           void *stream) {
  if (frames > STATIC_REFINE_MAX_FRAMES || n_mesh > STATIC_REFINE_MAX_MESH) {
    plr2::set_error("plr2_refine_loss_fwd_static: frames=%d / n_mesh=%d beyond the static scratch (max %d frames, %d mesh "
                    "points): use plr2_refine_loss_fwd with a caller-provided workspace",
                    frames, n_mesh, STATIC_REFINE_MAX_FRAMES, STATIC_REFINE_MAX_MESH);
    return PLR2_E_UNSUPPORTED;
  }
  if (frames <= 0 || n_mesh <= 0)      // shape errors / the empty batch: same answers as the workspace entry, no scratch needed
    return plr2_refine_loss_fwd(pred_r, pred_t, target, model_points, points, obj, sym_lut, num_obj, frames, n_points,
                                n_mesh, knn_mode, dis, new_points, new_target, jac_r, jac_t, pose_out, nullptr, 0, stream);
  void *scratch = nullptr;
  const int rc = refine_static_scratch_ptr(&scratch);
  if (rc) return rc;
  // plr2_refine_loss_fwd clears the arrival counters on `stream` before the launch (the array keeps the previous
  // call's contents; the partial sums are fully overwritten before they are read).  It allocates nothing.
  return plr2_refine_loss_fwd(pred_r, pred_t, target, model_points, points, obj, sym_lut, num_obj, frames, n_points,
                              n_mesh, knn_mode, dis, new_points, new_target, jac_r, jac_t, pose_out, scratch,
                              STATIC_REFINE_BYTES, stream);
}
