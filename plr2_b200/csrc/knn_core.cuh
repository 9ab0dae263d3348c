// Shared 1-NN scan over a shared-memory resident SoA reference tile (dim 3), used by the
// stand-alone KNN kernel and by the fused ADD-S loss kernels.
#pragma once
#include "common.cuh"
#include <math_constants.h>

namespace plr2 {

constexpr int KNN_CHUNK = 16;  // references per bookkeeping chunk (index recovery granularity)
#ifndef PLR2_KNN_UNROLL
#define PLR2_KNN_UNROLL 2
#endif
constexpr int KNN_SCAN_UNROLL = PLR2_KNN_UNROLL;   // chunks per unrolled iteration of the scan loop (tuning knob)

__host__ __device__ constexpr int knn_pad(int n) { return (n + KNN_CHUNK - 1) / KNN_CHUNK * KNN_CHUNK; }

// Distance used for ranking.
//   EXACT: q* hold -q                 -> fma(dz,dz,fma(dy,dy,dx*dx))  (knn.cu:88-91 after contraction)
//   FAST : q* hold -2(q - c), rw=|r-c|^2 -> |r-c|^2 - 2 (q-c).(r-c)
template <bool FAST>
__device__ __forceinline__ float pair_dist_scalar(float rx, float ry, float rz, float rw,
                                                  float qx, float qy, float qz) {
  if (FAST) {
    return fmaf(qx, rx, fmaf(qy, ry, fmaf(qz, rz, rw)));
  } else {
    const float dx = rx + qx, dy = ry + qy, dz = rz + qz;
    return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
  }
}

// Scan n_pad (multiple of KNN_CHUNK) references held in smem; two references per instruction via
// FADD2/FMUL2/FFMA2 (the query enters through the instruction's scalar-broadcast operand), minimum
// via FMNMX3.  The kernel is bound by the dispatch port the FMA and ALU pipes share (ncu:
// fma% + alu% = 95 %), so everything that is not an FFMA2 or an FMNMX3 is kept out of the loop:
//   * the four float4 rows of the NEXT group of 4 references are loaded while the current group
//     is evaluated (explicit register double buffer: with 8 queries per lane the compiler otherwise
//     issues the LDS right in front of their use and the warps sit on the short scoreboard);
//   * best[] itself is the running minimum; only the chunk id is selected per chunk
//     (strict '<' against the value at chunk entry => the EARLIEST chunk attaining the minimum).
template <bool FAST>
struct NnRefs4 {
  float4 X, Y, Z, W;
  __device__ __forceinline__ void load(const float *__restrict__ sx, const float *__restrict__ sy,
                                       const float *__restrict__ sz, const float *__restrict__ sw, int o) {
    X = *reinterpret_cast<const float4 *>(sx + o);
    Y = *reinterpret_cast<const float4 *>(sy + o);
    Z = *reinterpret_cast<const float4 *>(sz + o);
    if (FAST) W = *reinterpret_cast<const float4 *>(sw + o);
    else W = make_float4(0.f, 0.f, 0.f, 0.f);
  }
};

template <bool FAST, int QPT, bool PACKED = true>
__device__ __forceinline__ void nn_scan(const float *__restrict__ sx, const float *__restrict__ sy,
                                        const float *__restrict__ sz, const float *__restrict__ sw,
                                        int n_pad, int chunk_base, const float (&nqx)[QPT],
                                        const float (&nqy)[QPT], const float (&nqz)[QPT],
                                        float (&best)[QPT], int (&bchunk)[QPT]) {
  const int nchunk = n_pad / KNN_CHUNK;
  if (nchunk <= 0) return;
  NnRefs4<FAST> cur, nxt;
  cur.load(sx, sy, sz, sw, 0);
#pragma unroll KNN_SCAN_UNROLL
  for (int c = 0; c < nchunk; ++c) {
    float m[QPT];
#pragma unroll
    for (int i = 0; i < QPT; ++i) m[i] = best[i];
#pragma unroll
    for (int p = 0; p < KNN_CHUNK / 4; ++p) {
      // prefetch the next group (the last group of the scan re-reads itself)
      int on = c * KNN_CHUNK + p * 4 + 4;
      if (p == KNN_CHUNK / 4 - 1 && c + 1 >= nchunk) on -= 4;
      nxt.load(sx, sy, sz, sw, on);
      const float4 X = cur.X, Y = cur.Y, Z = cur.Z, W = cur.W;
      if (!PACKED) {
        // scalar form: same IEEE operations in the same order (bit-identical results), one
        // reference per instruction
#pragma unroll
        for (int i = 0; i < QPT; ++i) {
          const float a0 = pair_dist_scalar<FAST>(X.x, Y.x, Z.x, W.x, nqx[i], nqy[i], nqz[i]);
          const float a1 = pair_dist_scalar<FAST>(X.y, Y.y, Z.y, W.y, nqx[i], nqy[i], nqz[i]);
          const float a2 = pair_dist_scalar<FAST>(X.z, Y.z, Z.z, W.z, nqx[i], nqy[i], nqz[i]);
          const float a3 = pair_dist_scalar<FAST>(X.w, Y.w, Z.w, W.w, nqx[i], nqy[i], nqz[i]);
          m[i] = min3(min3(m[i], a0, a1), a2, a3);
        }
      } else {
        const f32x2 x01 = pack2(X.x, X.y), x23 = pack2(X.z, X.w);
        const f32x2 y01 = pack2(Y.x, Y.y), y23 = pack2(Y.z, Y.w);
        const f32x2 z01 = pack2(Z.x, Z.y), z23 = pack2(Z.z, Z.w);
        const f32x2 w01 = pack2(W.x, W.y), w23 = pack2(W.z, W.w);
#pragma unroll
        for (int i = 0; i < QPT; ++i) {
          const f32x2 qx = pack2(nqx[i], nqx[i]);
          const f32x2 qy = pack2(nqy[i], nqy[i]);
          const f32x2 qz = pack2(nqz[i], nqz[i]);
          f32x2 d01, d23;
          if (FAST) {
            d01 = fma2(qx, x01, fma2(qy, y01, fma2(qz, z01, w01)));
            d23 = fma2(qx, x23, fma2(qy, y23, fma2(qz, z23, w23)));
          } else {
            const f32x2 dx01 = add2(x01, qx), dx23 = add2(x23, qx);
            const f32x2 dy01 = add2(y01, qy), dy23 = add2(y23, qy);
            const f32x2 dz01 = add2(z01, qz), dz23 = add2(z23, qz);
            d01 = fma2(dz01, dz01, fma2(dy01, dy01, mul2(dx01, dx01)));
            d23 = fma2(dz23, dz23, fma2(dy23, dy23, mul2(dx23, dx23)));
          }
          float a0, a1, a2, a3;
          unpack2(d01, a0, a1);
          unpack2(d23, a2, a3);
          m[i] = min3(min3(m[i], a0, a1), a2, a3);
        }
      }
      cur = nxt;
    }
#pragma unroll
    for (int i = 0; i < QPT; ++i) {
      // m <= best always (m started from best; min ignores NaN), so best simply becomes m
      bchunk[i] = (m[i] < best[i]) ? chunk_base + c : bchunk[i];
      best[i] = m[i];
    }
  }
}

// Lowest index in chunk `chunk` whose ranking distance equals `best`, reading the (fully
// resident) smem tile.  Lanes walk the chunk rotated by lane id (different lanes generally hold
// different chunks), hence the min over matching indices.  Returns 0 when nothing matches
// (all distances +inf/NaN: the reference kernel keeps its initial index, knn.cu:127-128).
template <bool FAST>
__device__ __forceinline__ int nn_resolve_smem(const float *__restrict__ sx, const float *__restrict__ sy,
                                               const float *__restrict__ sz, const float *__restrict__ sw,
                                               int n_ref, int chunk, float best, float nqx, float nqy,
                                               float nqz, int lane) {
  const int base = chunk * KNN_CHUNK;
  int win = 0x7fffffff;
#pragma unroll 4
  for (int s = 0; s < KNN_CHUNK; ++s) {
    const int j = base + ((s + lane) & (KNN_CHUNK - 1));
    if (j < n_ref) {
      const float d = pair_dist_scalar<FAST>(sx[j], sy[j], sz[j], FAST ? sw[j] : 0.f, nqx, nqy, nqz);
      if (d == best) win = min(win, j);
    }
  }
  return win == 0x7fffffff ? 0 : win;
}

// Same contract as nn_resolve_smem, for callers whose lanes all hold different chunks anyway:
// the chunk is read as float4 rows and walked from its last reference to its first, so the last
// match written is the lowest index (no min, no rotation).
template <bool FAST>
__device__ __forceinline__ int nn_resolve_smem4(const float *__restrict__ sx, const float *__restrict__ sy,
                                                const float *__restrict__ sz, const float *__restrict__ sw,
                                                int n_ref, int chunk, float best, float nqx, float nqy,
                                                float nqz) {
  const int base = chunk * KNN_CHUNK;
  int win = -1;
#pragma unroll
  for (int p = KNN_CHUNK / 4 - 1; p >= 0; --p) {
    const int o = base + p * 4;
    const float4 X = *reinterpret_cast<const float4 *>(sx + o);
    const float4 Y = *reinterpret_cast<const float4 *>(sy + o);
    const float4 Z = *reinterpret_cast<const float4 *>(sz + o);
    float4 W = make_float4(0.f, 0.f, 0.f, 0.f);
    if (FAST) W = *reinterpret_cast<const float4 *>(sw + o);
    if (pair_dist_scalar<FAST>(X.w, Y.w, Z.w, W.w, nqx, nqy, nqz) == best) win = o + 3;
    if (pair_dist_scalar<FAST>(X.z, Y.z, Z.z, W.z, nqx, nqy, nqz) == best) win = o + 2;
    if (pair_dist_scalar<FAST>(X.y, Y.y, Z.y, W.y, nqx, nqy, nqz) == best) win = o + 1;
    if (pair_dist_scalar<FAST>(X.x, Y.x, Z.x, W.x, nqx, nqy, nqz) == best) win = o;
  }
  // a match can only be padding (index >= n_ref) when no real reference matched
  return (win < 0 || win >= n_ref) ? 0 : win;
}

}  // namespace plr2
