// Shared 1-NN scan over a shared-memory resident SoA reference tile (dim 3), used by the
// stand-alone KNN kernel and by the fused ADD-S loss kernels.
#pragma once
#include "common.cuh"
#include <math_constants.h>

namespace plr2 {

constexpr int KNN_CHUNK = 16;  // references per bookkeeping chunk (index recovery granularity)

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
// FADD2/FMUL2/FFMA2, minimum via FMNMX3.  best/bchunk are updated with strict '<' so the EARLIEST
// chunk attaining the running minimum is remembered.
template <bool FAST, int QPT>
__device__ __forceinline__ void nn_scan(const float *__restrict__ sx, const float *__restrict__ sy,
                                        const float *__restrict__ sz, const float *__restrict__ sw,
                                        int n_pad, int chunk_base, const float (&nqx)[QPT],
                                        const float (&nqy)[QPT], const float (&nqz)[QPT],
                                        float (&best)[QPT], int (&bchunk)[QPT]) {
  const int nchunk = n_pad / KNN_CHUNK;
  for (int c = 0; c < nchunk; ++c) {
    float m[QPT];
#pragma unroll
    for (int i = 0; i < QPT; ++i) m[i] = best[i];
#pragma unroll
    for (int p = 0; p < KNN_CHUNK / 4; ++p) {
      const int o = c * KNN_CHUNK + p * 4;
      const float4 X = *reinterpret_cast<const float4 *>(sx + o);
      const float4 Y = *reinterpret_cast<const float4 *>(sy + o);
      const float4 Z = *reinterpret_cast<const float4 *>(sz + o);
      const f32x2 x01 = pack2(X.x, X.y), x23 = pack2(X.z, X.w);
      const f32x2 y01 = pack2(Y.x, Y.y), y23 = pack2(Y.z, Y.w);
      const f32x2 z01 = pack2(Z.x, Z.y), z23 = pack2(Z.z, Z.w);
      f32x2 w01 = 0, w23 = 0;
      if (FAST) {
        const float4 W = *reinterpret_cast<const float4 *>(sw + o);
        w01 = pack2(W.x, W.y);
        w23 = pack2(W.z, W.w);
      }
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
#pragma unroll
    for (int i = 0; i < QPT; ++i) {
      if (m[i] < best[i]) {
        best[i] = m[i];
        bchunk[i] = chunk_base + c;
      }
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

}  // namespace plr2
