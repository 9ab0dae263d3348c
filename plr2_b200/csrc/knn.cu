// Kernel (4): brute-force nearest neighbour, replaces cuComputeDistanceGlobal + cuInsertionSort
// (reference lib/knn/src/cuda/knn.cu:36-101,113-176) without the ref_nb x query_nb scratch matrix.
//
// dim==3, k==1 (the only shape the losses use, lib/loss.py:43, lib/loss_refiner.py:43):
//   * references are staged once per CTA in shared memory as SoA (x[], y[], z[]) so that one
//     LDS.128 broadcast feeds 4 references to every lane;
//   * each thread keeps KNN_QPT queries in registers and evaluates TWO references per instruction
//     with the Blackwell packed FP32x2 ops (FADD2/FMUL2/FFMA2): per 2 pairs
//         EXACT: 3 FADD2 + FMUL2 + 2 FFMA2 + 1 FMNMX3      (bit-identical to knn.cu:88-91)
//         FAST : 3 FFMA2 + 1 FMNMX3 on centred coordinates  (|r|^2 - 2 q.r)
//   * only the running minimum is tracked in the hot loop; the winning 16-reference chunk is
//     remembered (strict '<' => earliest chunk) and re-scanned once at the end to recover the
//     lowest index attaining the minimum (knn.cu:158 semantics), written 1-based (knn.cu:171).
#include "knn_core.cuh"

namespace plr2 {

constexpr int KNN_THREADS = 128;
constexpr int KNN_QPT = 4;        // queries per thread
constexpr int KNN_TILE_R = 2048;  // references staged per shared-memory tile (3-4 x 8 KB)
constexpr int KNN_QPB = KNN_THREADS * KNN_QPT;

template <bool FAST>
__global__ void __launch_bounds__(KNN_THREADS)
knn3_kernel(const float *__restrict__ ref, const float *__restrict__ query,
            int64_t *__restrict__ idx, float *__restrict__ dist, int n_ref, int n_query) {
  __shared__ __align__(16) float sx[KNN_TILE_R];
  __shared__ __align__(16) float sy[KNN_TILE_R];
  __shared__ __align__(16) float sz[KNN_TILE_R];
  __shared__ __align__(16) float sw[FAST ? KNN_TILE_R : 4];

  const int b = blockIdx.y;
  const float *rb = ref + (size_t)b * 3 * n_ref;
  const float *qb = query + (size_t)b * 3 * n_query;
  const int tid = threadIdx.x;
  const int q0 = blockIdx.x * KNN_QPB + tid;

  // centre for FAST mode: the first reference point (any point near the cloud will do)
  float cx = 0.f, cy = 0.f, cz = 0.f;
  if (FAST) {
    cx = __ldg(rb);
    cy = __ldg(rb + n_ref);
    cz = __ldg(rb + 2 * n_ref);
  }

  float nqx[KNN_QPT], nqy[KNN_QPT], nqz[KNN_QPT];
  float best[KNN_QPT];
  int bchunk[KNN_QPT];
#pragma unroll
  for (int i = 0; i < KNN_QPT; ++i) {
    int q = q0 + i * KNN_THREADS;
    float x = 0.f, y = 0.f, z = 0.f;
    if (q < n_query) {
      x = __ldg(qb + q);
      y = __ldg(qb + n_query + q);
      z = __ldg(qb + 2 * n_query + q);
    }
    if (FAST) {
      nqx[i] = -2.0f * (x - cx);
      nqy[i] = -2.0f * (y - cy);
      nqz[i] = -2.0f * (z - cz);
    } else {
      nqx[i] = -x;
      nqy[i] = -y;
      nqz[i] = -z;
    }
    best[i] = CUDART_INF_F;
    bchunk[i] = 0;
  }

  for (int r0 = 0; r0 < n_ref; r0 += KNN_TILE_R) {
    const int nt = min(KNN_TILE_R, n_ref - r0);
    const int nt_pad = (nt + KNN_CHUNK - 1) / KNN_CHUNK * KNN_CHUNK;
    __syncthreads();
    for (int j = tid; j < nt_pad; j += KNN_THREADS) {
      float x, y, z;
      if (j < nt) {
        x = __ldg(rb + r0 + j);
        y = __ldg(rb + n_ref + r0 + j);
        z = __ldg(rb + 2 * n_ref + r0 + j);
        if (FAST) {
          x -= cx; y -= cy; z -= cz;
          sw[j] = fmaf(z, z, fmaf(y, y, x * x));
        }
      } else {
        // padding never wins: EXACT -> distance +inf (or NaN, ignored); FAST -> w = +inf
        x = y = z = FAST ? 0.f : CUDART_INF_F;
        if (FAST) sw[j] = CUDART_INF_F;
      }
      sx[j] = x; sy[j] = y; sz[j] = z;
    }
    __syncthreads();

    nn_scan<FAST, KNN_QPT>(sx, sy, sz, sw, nt_pad, r0 / KNN_CHUNK, nqx, nqy, nqz, best, bchunk);
  }

  // Recover the lowest reference index attaining best[i] by re-scanning the winning chunk from
  // global/L1 (the smem tile may hold another range by now).  Lanes walk the chunk rotated by
  // lane id, hence the min over matching indices.
  const int lane = tid & 31;
#pragma unroll
  for (int i = 0; i < KNN_QPT; ++i) {
    const int q = q0 + i * KNN_THREADS;
    if (q >= n_query) continue;
    const int base = bchunk[i] * KNN_CHUNK;
    int win = 0x7fffffff;
    for (int s = 0; s < KNN_CHUNK; ++s) {
      const int j = base + ((s + lane) & (KNN_CHUNK - 1));
      if (j < n_ref) {
        float rx = __ldg(rb + j), ry = __ldg(rb + n_ref + j), rz = __ldg(rb + 2 * n_ref + j);
        float rw = 0.f;
        if (FAST) {
          rx -= cx; ry -= cy; rz -= cz;
          rw = fmaf(rz, rz, fmaf(ry, ry, rx * rx));
        }
        const float d = pair_dist_scalar<FAST>(rx, ry, rz, rw, nqx[i], nqy[i], nqz[i]);
        if (d == best[i]) win = min(win, j);
      }
    }
    // No reference compared '<' +inf (all distances NaN/inf): the reference kernel keeps its
    // initial index 1 (knn.cu:127-128).
    if (win == 0x7fffffff) win = 0;
    if (!FAST) {
      // knn.cu:127: max_dist starts as dist[0]; a NaN there makes every later '<' false.
      const float d0 = pair_dist_scalar<false>(__ldg(rb), __ldg(rb + n_ref), __ldg(rb + 2 * n_ref),
                                               0.f, nqx[i], nqy[i], nqz[i]);
      if (d0 != d0) win = 0;
    }
    idx[(size_t)b * n_query + q] = (int64_t)win + 1;
    if (dist != nullptr) {
      // always report the direct-form squared distance on the original coordinates
      const float qx = -__ldg(qb + q), qy = -__ldg(qb + n_query + q), qz = -__ldg(qb + 2 * n_query + q);
      dist[(size_t)b * n_query + q] = pair_dist_scalar<false>(
          __ldg(rb + win), __ldg(rb + n_ref + win), __ldg(rb + 2 * n_ref + win), 0.f, qx, qy, qz);
    }
  }
}

// Generic (any dim, k <= 64): one thread per query, stable top-k by strict '<'
// (equivalent to the reference's insertion sort, knn.cu:131-173).
constexpr int KNN_KMAX = 64;
__global__ void __launch_bounds__(128)
knn_generic_kernel(const float *__restrict__ ref, const float *__restrict__ query,
                   int64_t *__restrict__ idx, float *__restrict__ dist, int dim, int n_ref,
                   int n_query, int k) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_query) return;
  const float *rb = ref + (size_t)b * dim * n_ref;
  const float *qb = query + (size_t)b * dim * n_query;
  float td[KNN_KMAX];
  int ti[KNN_KMAX];
  int cnt = 0;
  for (int j = 0; j < n_ref; ++j) {
    float ssd = 0.f;
    for (int d = 0; d < dim; ++d) {
      const float t = __ldg(rb + (size_t)d * n_ref + j) - __ldg(qb + (size_t)d * n_query + q);
      ssd = fmaf(t, t, ssd);
    }
    if (cnt < k) {
      int p = cnt++;
      while (p > 0 && td[p - 1] > ssd) { td[p] = td[p - 1]; ti[p] = ti[p - 1]; --p; }
      td[p] = ssd; ti[p] = j;
    } else if (ssd < td[k - 1]) {
      int p = k - 1;
      while (p > 0 && td[p - 1] > ssd) { td[p] = td[p - 1]; ti[p] = ti[p - 1]; --p; }
      td[p] = ssd; ti[p] = j;
    }
  }
  for (int i = 0; i < k; ++i) {
    // k > n_ref leaves the reference's rows untouched; we write index 0 there
    const size_t o = ((size_t)b * k + i) * n_query + q;
    idx[o] = i < cnt ? (int64_t)ti[i] + 1 : 0;
    if (dist != nullptr) dist[o] = i < cnt ? td[i] : CUDART_INF_F;
  }
}

}  // namespace plr2

extern "C" size_t plr2_knn_workspace_bytes(int, int, int, int, int) { return 0; }

extern "C" int plr2_knn_f32(const float *ref, const float *query, int64_t *idx, float *dist,
                            int batch, int dim, int n_ref, int n_query, int k, int mode,
                            void *, size_t, void *stream) {
  using namespace plr2;
  if (batch < 0 || dim <= 0 || n_ref <= 0 || n_query < 0 || k <= 0) {
    set_error("plr2_knn_f32: bad shape batch=%d dim=%d n_ref=%d n_query=%d k=%d", batch, dim, n_ref,
              n_query, k);
    return PLR2_E_SHAPE;
  }
  if (batch == 0 || n_query == 0) return 0;
  if (!ref || !query || !idx) { set_error("plr2_knn_f32: null pointer"); return PLR2_E_NULLPTR; }
  if (batch > 65535) { set_error("plr2_knn_f32: batch > 65535"); return PLR2_E_SHAPE; }
  cudaStream_t st = (cudaStream_t)stream;
  if (dim == 3 && k == 1) {
    dim3 grid((n_query + KNN_QPB - 1) / KNN_QPB, batch);
    if (mode == PLR2_KNN_FAST)
      knn3_kernel<true><<<grid, KNN_THREADS, 0, st>>>(ref, query, idx, dist, n_ref, n_query);
    else
      knn3_kernel<false><<<grid, KNN_THREADS, 0, st>>>(ref, query, idx, dist, n_ref, n_query);
  } else {
    if (k > KNN_KMAX) { set_error("plr2_knn_f32: k=%d > %d unsupported", k, KNN_KMAX); return PLR2_E_UNSUPPORTED; }
    dim3 grid((n_query + 127) / 128, batch);
    knn_generic_kernel<<<grid, 128, 0, st>>>(ref, query, idx, dist, dim, n_ref, n_query, k);
  }
  return check_launch("plr2_knn_f32");
}
