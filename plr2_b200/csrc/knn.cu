// Kernel (4): brute-force nearest neighbour, replaces cuComputeDistanceGlobal + cuInsertionSort
// (reference lib/knn/src/cuda/knn.cu:36-101,113-176) without the ref_nb x query_nb scratch matrix.
//
// dim==3, k==1 (the only shape the losses use, lib/loss.py:43, lib/loss_refiner.py:43):
//   * references are staged once per CTA in shared memory as SoA (x[], y[], z[], |r|^2[]; dynamic
//     smem sized to n_ref, tiles of KNN_TILE_MAX beyond that) so that one LDS.128 broadcast feeds
//     4 references to every lane;
//   * each lane keeps up to KNN_QPT queries in registers and evaluates TWO references per
//     instruction with the Blackwell packed FP32x2 ops (FADD2/FMUL2/FFMA2): per 2 pairs
//         EXACT: 3 FADD2 + FMUL2 + 2 FFMA2 + 1 FMNMX3      (bit-identical to knn.cu:88-91)
//         FAST : 3 FFMA2 + 1 FMNMX3 on centred coordinates  (|r|^2 - 2 q.r)
//   * only the running minimum is tracked in the hot loop; the winning 16-reference chunk is
//     remembered (strict '<' => earliest chunk) and re-read from smem once at the end to recover
//     the lowest index attaining the minimum (knn.cu:158 semantics), written 1-based (knn.cu:171).
//
// Work split (knn3_plan): grid = (nsplit, batch), CTA x owns queries [x*part, (x+1)*part) of its
// batch element and walks them in rounds of 128 lanes x QPT queries, QPT from {8,4,2,1} by binary
// decomposition of the per-lane count, so a ragged tail costs its own size and not a full round.
// Measured on B200 (tools/knn_split_sweep.py): the loop is bound by the dispatch port shared by the
// FMA and ALU pipes and needs >= 4 resident warps per scheduler to stay there.  ONE wave of long
// CTAs loses 12 % (the warp scheduler drains warps one after the other, the last ones run alone);
// several waves of short CTAs (<= 9 queries per lane) keep every scheduler full until the end.
#include "knn_core.cuh"

namespace plr2 {

// Tuning knobs (compile-time; the defaults are the shipped configuration, tools/knn_variants.sh sweeps them)
#ifndef PLR2_KNN_THREADS
#define PLR2_KNN_THREADS 128
#endif
#ifndef PLR2_KNN_MINB
#define PLR2_KNN_MINB 4
#endif
#ifndef PLR2_KNN_QPT
#define PLR2_KNN_QPT 8
#endif
constexpr int KNN_THREADS = PLR2_KNN_THREADS;
constexpr int KNN_QPT = PLR2_KNN_QPT;   // queries per lane in a full round
constexpr int KNN_TILE_MAX = 4096;  // references per smem tile (64 KB in FAST mode)

template <bool FAST>
__device__ __forceinline__ void knn3_stage(float *sx, float *sy, float *sz, float *sw,
                                           const float *__restrict__ rb, int n_ref, int r0, int nt,
                                           int nt_pad, float cx, float cy, float cz, int tid) {
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
      x = y = z = FAST ? 0.f : CUDART_INF_F;
      if (FAST) sw[j] = CUDART_INF_F;
    }
    sx[j] = x; sy[j] = y; sz[j] = z;
  }
}

template <bool FAST, bool PACKED, int QPT>
__device__ __forceinline__ void knn3_round(float *sx, float *sy, float *sz, float *sw,
                                           const float *__restrict__ rb, const float *__restrict__ qb,
                                           int64_t *__restrict__ idx_b, float *__restrict__ dist_b,
                                           int n_ref, int n_query, int tile_r, bool single, int q0,
                                           int q_end, float cx, float cy, float cz, int tid) {
  float nqx[QPT], nqy[QPT], nqz[QPT], best[QPT];
  int bchunk[QPT];
#pragma unroll
  for (int i = 0; i < QPT; ++i) {
    const int q = q0 + i * KNN_THREADS + tid;
    float x = 0.f, y = 0.f, z = 0.f;
    if (q < q_end) {
      x = __ldg(qb + q);
      y = __ldg(qb + n_query + q);
      z = __ldg(qb + 2 * n_query + q);
    }
    if (FAST) {
      nqx[i] = -2.0f * (x - cx); nqy[i] = -2.0f * (y - cy); nqz[i] = -2.0f * (z - cz);
    } else {
      nqx[i] = -x; nqy[i] = -y; nqz[i] = -z;
    }
    best[i] = CUDART_INF_F;
    bchunk[i] = 0;
  }
  if (single) {
    nn_scan<FAST, QPT, PACKED>(sx, sy, sz, sw, knn_pad(n_ref), 0, nqx, nqy, nqz, best, bchunk);
  } else {
    for (int r0 = 0; r0 < n_ref; r0 += tile_r) {
      const int nt = min(tile_r, n_ref - r0);
      __syncthreads();
      knn3_stage<FAST>(sx, sy, sz, sw, rb, n_ref, r0, nt, knn_pad(nt), cx, cy, cz, tid);
      __syncthreads();
      nn_scan<FAST, QPT, PACKED>(sx, sy, sz, sw, knn_pad(nt), r0 / KNN_CHUNK, nqx, nqy, nqz, best, bchunk);
    }
  }
  const int lane = tid & 31;
#pragma unroll
  for (int i = 0; i < QPT; ++i) {
    const int q = q0 + i * KNN_THREADS + tid;
    if (q >= q_end) continue;
    int win;
    if (single) {
      win = nn_resolve_smem4<FAST>(sx, sy, sz, sw, n_ref, bchunk[i], best[i], nqx[i], nqy[i], nqz[i]);
    } else {
      const int base = bchunk[i] * KNN_CHUNK;
      win = 0x7fffffff;
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
      if (win == 0x7fffffff) win = 0;
    }
    if (!FAST) {
      // knn.cu:127: max_dist starts as dist[0]; a NaN there makes every later '<' false.
      const float d0 = pair_dist_scalar<false>(__ldg(rb), __ldg(rb + n_ref), __ldg(rb + 2 * n_ref),
                                               0.f, nqx[i], nqy[i], nqz[i]);
      if (d0 != d0) win = 0;
    }
    idx_b[q] = (int64_t)win + 1;
    if (dist_b != nullptr) {
      const float qx = -__ldg(qb + q), qy = -__ldg(qb + n_query + q), qz = -__ldg(qb + 2 * n_query + q);
      dist_b[q] = pair_dist_scalar<false>(__ldg(rb + win), __ldg(rb + n_ref + win),
                                          __ldg(rb + 2 * n_ref + win), 0.f, qx, qy, qz);
    }
  }
}

template <bool FAST>
__global__ void __launch_bounds__(KNN_THREADS, PLR2_KNN_MINB)
knn3_kernel(const float *__restrict__ ref, const float *__restrict__ query,
                     int64_t *__restrict__ idx, float *__restrict__ dist, int n_ref, int n_query,
                     int part, int tile_r) {
  extern __shared__ __align__(16) float knn_smem[];
  float *sx = knn_smem, *sy = sx + tile_r, *sz = sy + tile_r, *sw = sz + tile_r;
  const int b = blockIdx.y, tid = threadIdx.x;
  const float *rb = ref + (size_t)b * 3 * n_ref;
  const float *qb = query + (size_t)b * 3 * n_query;
  int64_t *idx_b = idx + (size_t)b * n_query;
  float *dist_b = dist ? dist + (size_t)b * n_query : nullptr;
  const int q_beg = blockIdx.x * part;
  const int q_end = min(n_query, q_beg + part);
  if (q_beg >= q_end) return;

  float cx = 0.f, cy = 0.f, cz = 0.f;
  if (FAST) { cx = __ldg(rb); cy = __ldg(rb + n_ref); cz = __ldg(rb + 2 * n_ref); }
  const bool single = n_ref <= tile_r;
  if (single) {
    knn3_stage<FAST>(sx, sy, sz, sw, rb, n_ref, 0, n_ref, knn_pad(n_ref), cx, cy, cz, tid);
    __syncthreads();
  }
  constexpr int QMAX = KNN_QPT;
  int q = q_beg;
  while (q < q_end) {
    const int rem = (q_end - q + KNN_THREADS - 1) / KNN_THREADS;  // queries left per lane
    if (rem >= QMAX) {
      knn3_round<FAST, true, QMAX>(sx, sy, sz, sw, rb, qb, idx_b, dist_b, n_ref, n_query, tile_r, single, q, q_end, cx, cy, cz, tid);
      q += QMAX * KNN_THREADS;
    } else if (QMAX > 4 && rem >= 4) {
      knn3_round<FAST, true, 4>(sx, sy, sz, sw, rb, qb, idx_b, dist_b, n_ref, n_query, tile_r, single, q, q_end, cx, cy, cz, tid);
      q += 4 * KNN_THREADS;
    } else if (rem >= 2) {
      knn3_round<FAST, true, 2>(sx, sy, sz, sw, rb, qb, idx_b, dist_b, n_ref, n_query, tile_r, single, q, q_end, cx, cy, cz, tid);
      q += 2 * KNN_THREADS;
    } else {
      knn3_round<FAST, true, 1>(sx, sy, sz, sw, rb, qb, idx_b, dist_b, n_ref, n_query, tile_r, single, q, q_end, cx, cy, cz, tid);
      q += KNN_THREADS;
    }
  }
}

// Host-side plan: split each batch element's queries into equal parts of `per` queries per lane.
// Cost model (relative units fitted to the sweep): a round of 8/4/2/1 queries per lane costs
// 8/4.6/2.8/1.9, a CTA costs its rounds + 0.35 (staging, launch), the grid costs
// ceil(CTAs / resident slots) CTA times.  Parts longer than 9 queries per lane are not considered
// (see the header: long CTAs drain badly) unless the batch element has fewer queries anyway.
struct Knn3Plan { int nsplit, part; };
static Knn3Plan knn3_plan(int batch, int n_query, int slots) {
  const int lanes_q = (n_query + KNN_THREADS - 1) / KNN_THREADS;
  Knn3Plan best{1, lanes_q * KNN_THREADS};
  double best_cost = 1e300;
  for (int per = 1; per <= 9 && per <= lanes_q; ++per) {
    const int nsplit = (lanes_q + per - 1) / per;
    const int per_eff = (lanes_q + nsplit - 1) / nsplit;  // evened out
    const double round_cost = 8.0 * (per_eff >> 3) + ((per_eff & 4) ? 4.6 : 0.0) +
                              ((per_eff & 2) ? 2.8 : 0.0) + ((per_eff & 1) ? 1.9 : 0.0);
    const long long ctas = (long long)nsplit * batch;
    const long long waves = (ctas + slots - 1) / slots;
    const double cost = (double)waves * (round_cost + 0.35);
    if (cost < best_cost - 1e-9) { best_cost = cost; best = {nsplit, per_eff * KNN_THREADS}; }
  }
  best.nsplit = (n_query + best.part - 1) / best.part;
  return best;
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

extern "C" int plr2_knn_plan(int batch, int n_query, int resident_ctas, int *nsplit, int *part) {
  if (batch <= 0 || n_query <= 0 || resident_ctas <= 0 || !nsplit || !part) {
    plr2::set_error("plr2_knn_plan: bad arguments");
    return PLR2_E_SHAPE;
  }
  const plr2::Knn3Plan p = plr2::knn3_plan(batch, n_query, resident_ctas);
  *nsplit = p.nsplit;
  *part = p.part;
  return 0;
}

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
    const bool fast = mode == PLR2_KNN_FAST;
    const int tile_r = min(knn_pad(n_ref), KNN_TILE_MAX);
    const size_t smem = (size_t)tile_r * sizeof(float) * (fast ? 4 : 3);
    auto kern = fast ? knn3_kernel<true> : knn3_kernel<false>;
    static PerDeviceInt slots[2];  // resident CTAs per SM of the device, per mode (smem: n_ref <= 1024)
    static PerDeviceInt attr_set;  // the shared-memory attribute is per device
    const int dev = current_device();
    const int sms = device_sm_count();
    if (!attr_set.get(dev)) {
      cudaError_t e = cudaFuncSetAttribute(knn3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, KNN_TILE_MAX * 16);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(knn3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, KNN_TILE_MAX * 12);
      if (e != cudaSuccess) { set_error("plr2_knn_f32: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
      attr_set.set(dev, 1);
    }
    int per_sm = slots[fast].get(dev);
    if (!per_sm || tile_r > 1024) {
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, KNN_THREADS, smem);
      if (per_sm < 1) per_sm = 1;
      if (tile_r <= 1024) slots[fast].set(dev, per_sm);
    }
    const Knn3Plan plan = knn3_plan(batch, n_query, sms * per_sm);
    dim3 grid(plan.nsplit, batch);
    kern<<<grid, KNN_THREADS, smem, st>>>(ref, query, idx, dist, n_ref, n_query, plan.part, tile_r);
  } else {
    if (k > KNN_KMAX) { set_error("plr2_knn_f32: k=%d > %d unsupported", k, KNN_KMAX); return PLR2_E_UNSUPPORTED; }
    dim3 grid((n_query + 127) / 128, batch);
    knn_generic_kernel<<<grid, 128, 0, st>>>(ref, query, idx, dist, dim, n_ref, n_query, k);
  }
  return check_launch("plr2_knn_f32");
}
