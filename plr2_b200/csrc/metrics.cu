// Evaluation metrics after the path, on the device so that poses never leave it:
//   * ADD / ADD-S between two poses applied to the model points (src/estimation/errors.py:53-84),
//     the ADD-S nearest neighbour through the same shared-memory 1-NN scan as kernel (4);
//   * the YCB-Video accuracy-threshold AUC (replace_ycb_toolbox/plot_accuracy_keyframe.m:38-52,150-170)
//     without a sort: rank and predecessor of every distinct distance by brute force over all pairs.
#include "knn_core.cuh"

namespace plr2 {

constexpr int PM_THREADS = 128;
constexpr int PM_SLICES = 8;      // at most this many CTAs per frame, each owning a contiguous range of query points
constexpr int PM_QPT = 4;         // queries per thread and pass (as kernel (4))

// fixed-order block sum: warp shuffles, then warp 0 adds the per-warp values in warp order
template <typename T>
__device__ __forceinline__ T block_sum_ordered(T v, T *scratch) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  T r = 0;
  for (int w = 0; w < nw; ++w) r += scratch[w];
  return r;
}

// grid (PM_SLICES, frames); partial[(f*PM_SLICES + s)*2 + {0,1}] = sums of ADD / ADD-S distances of slice s
template <bool ADDS>
__global__ void __launch_bounds__(PM_THREADS)
pose_metric_kernel(const float *__restrict__ model, long long model_stride, const float *__restrict__ pose_a,
                   const float *__restrict__ pose_b, int M, float *__restrict__ partial) {
  extern __shared__ __align__(16) float sm[];
  __shared__ float scratch[PM_THREADS / 32];
  const int f = blockIdx.y, s = blockIdx.x, tid = threadIdx.x;
  const float *mp = model + (size_t)f * model_stride;
  float Ra[12], Rb[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) { Ra[i] = __ldg(pose_a + (size_t)f * 12 + i); Rb[i] = __ldg(pose_b + (size_t)f * 12 + i); }
  const int n_pad = knn_pad(M);
  float *sx = sm, *sy = sm + n_pad, *sz = sm + 2 * n_pad;
  if (ADDS) {                                   // model_1 = pose_a applied: the reference set of the 1-NN
    for (int j = tid; j < n_pad; j += PM_THREADS) {
      float x = CUDART_INF_F, y = CUDART_INF_F, z = CUDART_INF_F;     // padding never wins
      if (j < M) {
        const float px = __ldg(mp + 3 * j), py = __ldg(mp + 3 * j + 1), pz = __ldg(mp + 3 * j + 2);
        x = fmaf(Ra[0], px, fmaf(Ra[1], py, fmaf(Ra[2], pz, Ra[9])));
        y = fmaf(Ra[3], px, fmaf(Ra[4], py, fmaf(Ra[5], pz, Ra[10])));
        z = fmaf(Ra[6], px, fmaf(Ra[7], py, fmaf(Ra[8], pz, Ra[11])));
      }
      sx[j] = x; sy[j] = y; sz[j] = z;
    }
    __syncthreads();
  }
  const int per = (M + gridDim.x - 1) / gridDim.x;
  const int j0 = s * per, j1 = min(M, j0 + per);
  float acc_add = 0.f, acc_adds = 0.f;
  for (int base = j0; base < j1; base += PM_THREADS * PM_QPT) {
    float nqx[PM_QPT], nqy[PM_QPT], nqz[PM_QPT], best[PM_QPT];
    int bchunk[PM_QPT];
    bool live[PM_QPT];
#pragma unroll
    for (int i = 0; i < PM_QPT; ++i) {
      const int j = base + i * PM_THREADS + tid;
      live[i] = j < j1;
      float px = 0.f, py = 0.f, pz = 0.f;
      if (live[i]) { px = __ldg(mp + 3 * j); py = __ldg(mp + 3 * j + 1); pz = __ldg(mp + 3 * j + 2); }
      const float bx = fmaf(Rb[0], px, fmaf(Rb[1], py, fmaf(Rb[2], pz, Rb[9])));
      const float by = fmaf(Rb[3], px, fmaf(Rb[4], py, fmaf(Rb[5], pz, Rb[10])));
      const float bz = fmaf(Rb[6], px, fmaf(Rb[7], py, fmaf(Rb[8], pz, Rb[11])));
      const float ax = fmaf(Ra[0], px, fmaf(Ra[1], py, fmaf(Ra[2], pz, Ra[9])));
      const float ay = fmaf(Ra[3], px, fmaf(Ra[4], py, fmaf(Ra[5], pz, Ra[10])));
      const float az = fmaf(Ra[6], px, fmaf(Ra[7], py, fmaf(Ra[8], pz, Ra[11])));
      const float dx = ax - bx, dy = ay - by, dz = az - bz;
      if (live[i]) acc_add += sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
      nqx[i] = -bx; nqy[i] = -by; nqz[i] = -bz;
      best[i] = CUDART_INF_F; bchunk[i] = 0;
    }
    if (ADDS) {
      nn_scan<false, PM_QPT>(sx, sy, sz, nullptr, n_pad, 0, nqx, nqy, nqz, best, bchunk);
#pragma unroll
      for (int i = 0; i < PM_QPT; ++i)
        if (live[i]) acc_adds += sqrtf(best[i]);
    }
  }
  const float sum_add = block_sum_ordered(acc_add, scratch);
  const float sum_adds = ADDS ? block_sum_ordered(acc_adds, scratch) : 0.f;
  if (tid == 0) {
    partial[((size_t)f * PM_SLICES + s) * 2] = sum_add;
    partial[((size_t)f * PM_SLICES + s) * 2 + 1] = sum_adds;
  }
}

__global__ void pose_metric_finish_kernel(const float *__restrict__ partial, int frames, int M, int slices,
                                          float *__restrict__ add, float *__restrict__ adds) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= frames) return;
  float a = 0.f, b = 0.f;
  for (int s = 0; s < slices; ++s) {             // slice order: deterministic
    a += partial[((size_t)f * PM_SLICES + s) * 2];
    b += partial[((size_t)f * PM_SLICES + s) * 2 + 1];
  }
  if (add) add[f] = a / (float)M;
  if (adds) adds[f] = b / (float)M;
}

__global__ void pose_errors_kernel(const float *__restrict__ pa, const float *__restrict__ pb, int frames,
                                   float *__restrict__ rot_err, float *__restrict__ trans_err) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= frames) return;
  const float *A = pa + (size_t)f * 12, *B = pb + (size_t)f * 12;
  float D[9];                                              // D = Ra Rb^T
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      D[3 * i + j] = fmaf(A[3 * i], B[3 * j], fmaf(A[3 * i + 1], B[3 * j + 1], A[3 * i + 2] * B[3 * j + 2]));
  if (rot_err) {
    const float vx = D[7] - D[5], vy = D[2] - D[6], vz = D[3] - D[1];        // 2 sin(angle) * axis
    const float s2 = sqrtf(fmaf(vz, vz, fmaf(vy, vy, vx * vx)));
    const float c2 = D[0] + D[4] + D[8] - 1.0f;                               // 2 cos(angle)
    rot_err[f] = atan2f(s2, c2);
  }
  if (trans_err) {
    const float dx = A[9] - B[9], dy = A[10] - B[10], dz = A[11] - B[11];
    trans_err[f] = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
  }
}

// ---- AUC -----------------------------------------------------------------------------------
constexpr int AUC_THREADS = 256;
struct AucPartial { double contrib; double dmax; long long below; long long nfin; };

// A distance takes part ("finite" after plot_accuracy_keyframe.m:40 `D(D > max_distance) = inf` and VOCap's
// isfinite filter) iff it is not NaN, not +-inf and not above max_distance.
__device__ __forceinline__ float auc_key(float d, double maxd) {
  return (d == d && fabsf(d) != CUDART_INF_F && !((double)d > maxd)) ? d : CUDART_INF_F;
}

__global__ void __launch_bounds__(AUC_THREADS)
auc_partial_kernel(const float *__restrict__ dist, int n, double maxd, double below, AucPartial *__restrict__ partial) {
  __shared__ float tile[AUC_THREADS];
  __shared__ double sd[AUC_THREADS / 32];
  __shared__ long long sl[AUC_THREADS / 32];
  const int i = blockIdx.x * AUC_THREADS + threadIdx.x;
  const float di = i < n ? auc_key(__ldg(dist + i), maxd) : CUDART_INF_F;
  const bool fin = di != CUDART_INF_F;
  int cnt = 0;                 // # participating distances strictly below di
  float pred = 0.f;            // largest of them (mrec(1) = 0 when there is none)
  bool dup = false;            // an equal value sits at a lower index: only the first of a run contributes
  for (int j0 = 0; j0 < n; j0 += AUC_THREADS) {
    __syncthreads();
    const int j = j0 + threadIdx.x;
    tile[threadIdx.x] = j < n ? auc_key(__ldg(dist + j), maxd) : CUDART_INF_F;
    __syncthreads();
    const int lim = min(AUC_THREADS, n - j0);
    for (int t = 0; t < lim; ++t) {
      const float v = tile[t];
      if (v < di) { ++cnt; pred = fmaxf(pred, v); }
      dup = dup || (v == di && j0 + t < i);
    }
  }
  // mrec(i) - mrec(i-1) times mpre(i) = accuracy at the first element of the run (VOCap :165-170)
  const double c = (fin && !dup) ? ((double)di - (double)pred) * ((double)(cnt + 1) / (double)n) : 0.0;
  const double csum = block_sum_ordered<double>(c, sd);
  const long long nb = block_sum_ordered<long long>((fin && (double)di < below) ? 1 : 0, sl);
  const long long nf = block_sum_ordered<long long>(fin ? 1 : 0, sl);
  // block max of the participating distances
  float m = fin ? di : -CUDART_INF_F;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) tile[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < AUC_THREADS / 32; ++w) m = fmaxf(m, tile[w]);
    AucPartial p; p.contrib = csum; p.dmax = (double)m; p.below = nb; p.nfin = nf;
    partial[blockIdx.x] = p;
  }
}

__global__ void auc_finish_kernel(const AucPartial *__restrict__ partial, int blocks, int n, double maxd,
                                  double *__restrict__ out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double sum = 0.0, dmax = -1.0;
  long long below = 0, nfin = 0;
  for (int b = 0; b < blocks; ++b) {
    sum += partial[b].contrib; below += partial[b].below; nfin += partial[b].nfin;
    dmax = partial[b].dmax > dmax ? partial[b].dmax : dmax;
  }
  double ap = 0.0;
  if (n > 0 && nfin > 0) {
    sum += (maxd - dmax) * ((double)nfin / (double)n);      // mrec = [...; 0.1], mpre = [...; prec(end)]
    ap = sum * (1.0 / maxd);                               // `* 10` for max_distance 0.1 (VOCap :170)
  }
  out[0] = ap;
  out[1] = n > 0 ? (double)below / (double)n : 0.0;
}

}  // namespace plr2

using namespace plr2;

extern "C" size_t plr2_pose_metric_workspace_bytes(int frames) {
  return frames > 0 ? (size_t)frames * PM_SLICES * 2 * sizeof(float) : 0;
}

extern "C" int plr2_pose_metric(const float *model_points, long long model_stride, const float *pose_a,
                                const float *pose_b, int frames, int n_mesh, float *add, float *adds,
                                void *workspace, size_t workspace_bytes, void *stream) {
  if (frames < 0 || n_mesh <= 0 || frames > 65535 || model_stride < 0) {
    set_error("plr2_pose_metric: bad shape F=%d M=%d stride=%lld", frames, n_mesh, model_stride);
    return PLR2_E_SHAPE;
  }
  if (frames == 0) return 0;
  if (!model_points || !pose_a || !pose_b || (!add && !adds) || !workspace) {
    set_error("plr2_pose_metric: null pointer");
    return PLR2_E_NULLPTR;
  }
  if (n_mesh > 8192) { set_error("plr2_pose_metric: n_mesh <= 8192 (got %d)", n_mesh); return PLR2_E_UNSUPPORTED; }
  if (workspace_bytes < plr2_pose_metric_workspace_bytes(frames)) {
    set_error("plr2_pose_metric: workspace %zu < %zu bytes", workspace_bytes, plr2_pose_metric_workspace_bytes(frames));
    return PLR2_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float *partial = (float *)workspace;
  // slices sized to one pass of a CTA (PM_THREADS * PM_QPT queries), so no thread idles through the scan
  int slices = (n_mesh + PM_THREADS * PM_QPT - 1) / (PM_THREADS * PM_QPT);
  slices = slices > PM_SLICES ? PM_SLICES : slices;
  dim3 grid(slices, frames);
  if (adds) {
    const size_t smem = (size_t)knn_pad(n_mesh) * 3 * sizeof(float);
    static PerDeviceInt attr;     // per-device function attribute
    const int dev = current_device();
    if (!attr.get(dev)) {
      cudaError_t e = cudaFuncSetAttribute(pose_metric_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 3 * 4);
      if (e != cudaSuccess) { set_error("plr2_pose_metric: %s", cudaGetErrorString(e)); return (int)e; }
      attr.set(dev, 1);
    }
    pose_metric_kernel<true><<<grid, PM_THREADS, smem, st>>>(model_points, model_stride, pose_a, pose_b, n_mesh, partial);
  } else {
    pose_metric_kernel<false><<<grid, PM_THREADS, 0, st>>>(model_points, model_stride, pose_a, pose_b, n_mesh, partial);
  }
  pose_metric_finish_kernel<<<(frames + 127) / 128, 128, 0, st>>>(partial, frames, n_mesh, slices, add, adds);
  return check_launch("plr2_pose_metric", 2);
}

extern "C" int plr2_pose_errors(const float *pose_a, const float *pose_b, int frames, float *rot_err, float *trans_err,
                                void *stream) {
  if (frames < 0) { set_error("plr2_pose_errors: bad shape F=%d", frames); return PLR2_E_SHAPE; }
  if (frames == 0) return 0;
  if (!pose_a || !pose_b || (!rot_err && !trans_err)) { set_error("plr2_pose_errors: null pointer"); return PLR2_E_NULLPTR; }
  pose_errors_kernel<<<(frames + 127) / 128, 128, 0, (cudaStream_t)stream>>>(pose_a, pose_b, frames, rot_err, trans_err);
  return check_launch("plr2_pose_errors");
}

extern "C" size_t plr2_auc_workspace_bytes(int n) {
  const int blocks = n > 0 ? (n + AUC_THREADS - 1) / AUC_THREADS : 1;
  return (size_t)blocks * sizeof(AucPartial);
}

extern "C" int plr2_auc_f32(const float *dist, int n, double max_distance, double below, double *out,
                            void *workspace, size_t workspace_bytes, void *stream) {
  if (n < 0 || !(max_distance > 0.0)) {
    set_error("plr2_auc_f32: bad arguments n=%d max_distance=%g", n, max_distance);
    return PLR2_E_SHAPE;
  }
  if (!out || (n > 0 && (!dist || !workspace))) { set_error("plr2_auc_f32: null pointer"); return PLR2_E_NULLPTR; }
  if (n > 0 && workspace_bytes < plr2_auc_workspace_bytes(n)) {
    set_error("plr2_auc_f32: workspace %zu < %zu bytes", workspace_bytes, plr2_auc_workspace_bytes(n));
    return PLR2_E_WORKSPACE;
  }
  if (((uintptr_t)out & 7) || ((uintptr_t)workspace & 7)) { set_error("plr2_auc_f32: out/workspace must be 8-byte aligned"); return PLR2_E_ALIGN; }
  cudaStream_t st = (cudaStream_t)stream;
  const int blocks = n > 0 ? (n + AUC_THREADS - 1) / AUC_THREADS : 0;
  if (blocks) auc_partial_kernel<<<blocks, AUC_THREADS, 0, st>>>(dist, n, max_distance, below, (AucPartial *)workspace);
  auc_finish_kernel<<<1, 32, 0, st>>>((const AucPartial *)workspace, blocks, n, max_distance, out);
  return check_launch("plr2_auc_f32", blocks ? 2 : 1);
}
