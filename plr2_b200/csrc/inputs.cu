// Input step before the path: the per-frame numpy work of datasets/ycb/dataset.py:177-197,
// 215-219 (= src/loaders_v2/dataset_ycb.py:91-119) for F frames on the device:
//   choose_kernel          mask crop -> ascending non-zero list -> exactly n_points indices (wrap pad or subset)
//   backproject_kernel     depth + choose + intrinsics -> point cloud, float32 operation by operation as numpy
//   transform_target_kernel  model_points . R^T + t (+ add_t) in float64, rounded once
#include "common.cuh"

namespace plr2 {

constexpr int CH_THREADS = 1024;

// Block-wide exclusive scan of one int per thread (CH_THREADS threads); returns the exclusive prefix and the
// block total through `total`.  warp_tot is a [32] smem scratch.
__device__ __forceinline__ int block_excl_scan(int v, int *warp_tot, int &total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();                       // previous use of warp_tot is over
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = warp_tot[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    warp_tot[lane] = winc - w;           // exclusive prefix of the warp totals
    if (lane == 31) warp_tot[32] = winc; // block total
  }
  __syncthreads();
  total = warp_tot[32];
  return warp_tot[warp] + inc - v;
}

// 64 -> 32 bit mixing of (seed, frame, k): splitmix64 finaliser
__device__ __forceinline__ uint32_t choose_key(unsigned long long seed, int f, int k) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ull * ((unsigned long long)(unsigned)f * 0x100000001B3ull + (unsigned)k + 1ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return (uint32_t)((z ^ (z >> 31)) >> 32);
}

// one CTA per frame; list = workspace row of this frame (crop-flat indices of the non-zero pixels, ascending)
__global__ void __launch_bounds__(CH_THREADS)
choose_kernel(const uint8_t *__restrict__ mask, int img_h, int img_w, const int32_t *__restrict__ bbox, int N,
              const uint8_t *__restrict__ keep, long long keep_stride, unsigned long long seed,
              int64_t *__restrict__ choose, int32_t *__restrict__ count_out, int32_t *__restrict__ lists, int list_stride) {
  __shared__ int warp_tot[33];
  __shared__ int hist[256];
  __shared__ uint32_t s_prefix;
  __shared__ int s_remaining;
  const int f = blockIdx.x, tid = threadIdx.x;
  const int rmin = bbox[4 * f], rmax = bbox[4 * f + 1], cmin = bbox[4 * f + 2], cmax = bbox[4 * f + 3];
  const int ch = rmax - rmin, cw = cmax - cmin;
  const int P = (ch > 0 && cw > 0) ? ch * cw : 0;
  const uint8_t *img = mask + (size_t)f * img_h * img_w;
  int32_t *list = lists + (size_t)f * list_stride;
  int64_t *out = choose + (size_t)f * N;

  // ---- 1. ascending list of the non-zero crop pixels
  int K = 0;
  for (int p0 = 0; p0 < P; p0 += CH_THREADS) {
    const int p = p0 + tid;
    int nz = 0;
    if (p < P) {
      const int r = rmin + p / cw, c = cmin + p % cw;
      nz = (r >= 0 && r < img_h && c >= 0 && c < img_w && img[(size_t)r * img_w + c] != 0) ? 1 : 0;
    }
    int tot;
    const int pos = block_excl_scan(nz, warp_tot, tot);
    if (nz) list[K + pos] = p;
    K += tot;
  }
  __syncthreads();
  if (tid == 0 && count_out) count_out[f] = K;
  if (K == 0) {
    for (int i = tid; i < N; i += CH_THREADS) out[i] = 0;
    return;
  }
  // ---- 2a. few pixels: np.pad(choose, (0, N - K), 'wrap')
  if (K <= N) {
    for (int i = tid; i < N; i += CH_THREADS) out[i] = list[i % K];
    return;
  }
  // ---- 2b. subset of exactly N, ascending order kept
  uint32_t T = 0;        // random mode: the N-th smallest key
  int need_eq = 0;       //              how many keys == T are still taken (lowest k first)
  if (keep == nullptr) {
    // MSB-first radix select of the N-th smallest 32-bit key
    if (tid == 0) { s_prefix = 0; s_remaining = N; }
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      if (tid < 256) hist[tid] = 0;
      __syncthreads();
      const uint32_t prefix = s_prefix;
      const uint32_t himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      for (int k = tid; k < K; k += CH_THREADS) {
        const uint32_t key = choose_key(seed, f, k);
        if ((key & himask) == prefix) atomicAdd(&hist[(key >> shift) & 255], 1);
      }
      __syncthreads();
      if (tid == 0) {
        int rem = s_remaining, b = 0;
        while (b < 255 && hist[b] < rem) { rem -= hist[b]; ++b; }
        s_prefix = prefix | ((uint32_t)b << shift);
        s_remaining = rem;              // rank of the wanted key among those sharing the new prefix
      }
      __syncthreads();
    }
    T = s_prefix;
    need_eq = s_remaining;              // of the keys equal to T, the first need_eq (by k) are taken
  }
  const uint8_t *kp = keep ? keep + (size_t)f * keep_stride : nullptr;
  int written = 0, eq_seen = 0;
  for (int k0 = 0; k0 < K; k0 += CH_THREADS) {
    const int k = k0 + tid;
    int take = 0, eq = 0;
    if (k < K) {
      if (kp) take = kp[k] != 0;
      else {
        const uint32_t key = choose_key(seed, f, k);
        take = key < T;
        eq = key == T;
      }
    }
    int tot_eq = 0;
    if (!kp) {
      const int eq_pos = block_excl_scan(eq, warp_tot, tot_eq);
      if (eq && eq_seen + eq_pos < need_eq) take = 1;
    }
    int tot;
    const int pos = block_excl_scan(take, warp_tot, tot);
    if (take && written + pos < N) out[written + pos] = list[k];
    written += tot;
    eq_seen += tot_eq;
  }
  // a keep mask with fewer than N ones: the tail wraps over what was written (never happens for the
  // reference's c_mask, which holds exactly N ones)
  if (written < N) {
    __syncthreads();
    if (written == 0) { for (int i = tid; i < N; i += CH_THREADS) out[i] = 0; }
    else for (int i = written + tid; i < N; i += CH_THREADS) out[i] = out[i % written];
  }
}

template <typename DT>
__global__ void backproject_kernel(const DT *__restrict__ depth, int img_h, int img_w, const int32_t *__restrict__ bbox,
                                   const int64_t *__restrict__ choose, const float *__restrict__ cam,
                                   const double *__restrict__ add_t, int N, float *__restrict__ cloud) {
  const int f = blockIdx.y, n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int rmin = bbox[4 * f], cmin = bbox[4 * f + 2], cw = bbox[4 * f + 3] - cmin;
  const float cx = cam[5 * f], cy = cam[5 * f + 1], fx = cam[5 * f + 2], fy = cam[5 * f + 3], scale = cam[5 * f + 4];
  const long long c = choose[(size_t)f * N + n];
  const int r = rmin + (int)(c / (cw > 0 ? cw : 1)), col = cmin + (int)(c % (cw > 0 ? cw : 1));
  float d = 0.f;
  if (r >= 0 && r < img_h && col >= 0 && col < img_w) d = (float)depth[((size_t)f * img_h + r) * img_w + col];
  // float32 step by step (no contraction): pt2 = depth / cam_scale; pt0 = (ymap - cx) * pt2 / fx; ...
  const float z = __fdiv_rn(d, scale);
  float x = __fdiv_rn(__fmul_rn(__fsub_rn((float)col, cx), z), fx);
  float y = __fdiv_rn(__fmul_rn(__fsub_rn((float)r, cy), z), fy);
  float zz = z;
  if (add_t) {                                 // np.add(cloud, add_t) is a float64 add, rounded at .astype(float32)
    x = (float)((double)x + add_t[3 * f]);
    y = (float)((double)y + add_t[3 * f + 1]);
    zz = (float)((double)z + add_t[3 * f + 2]);
  }
  float *o = cloud + ((size_t)f * N + n) * 3;
  o[0] = x; o[1] = y; o[2] = zz;
}

__global__ void transform_target_kernel(const float *__restrict__ model, const double *__restrict__ R,
                                        const double *__restrict__ t, const double *__restrict__ add_t, int M,
                                        float *__restrict__ target) {
  const int f = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  const double *Rf = R + 9 * (size_t)f;
  const float *p = model + ((size_t)f * M + j) * 3;
  const double px = p[0], py = p[1], pz = p[2];
  float *o = target + ((size_t)f * M + j) * 3;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    // row j of model_points . R^T: sum_k p_k R[i][k], then + (t (+ add_t)) as np.add(target, target_t + add_t)
    double v = __dadd_rn(__dadd_rn(__dmul_rn(px, Rf[3 * i]), __dmul_rn(py, Rf[3 * i + 1])), __dmul_rn(pz, Rf[3 * i + 2]));
    const double off = add_t ? __dadd_rn(t[3 * f + i], add_t[3 * f + i]) : t[3 * f + i];
    o[i] = (float)__dadd_rn(v, off);
  }
}

}  // namespace plr2

using namespace plr2;

extern "C" size_t plr2_choose_workspace_bytes(int frames, int max_crop_pixels) {
  if (frames <= 0 || max_crop_pixels <= 0) return 0;
  return (size_t)frames * (size_t)max_crop_pixels * sizeof(int32_t);
}

extern "C" int plr2_choose_from_mask(const uint8_t *mask, int img_h, int img_w, const int32_t *bbox, int frames,
                                     int n_points, const uint8_t *keep, long long keep_stride,
                                     unsigned long long seed, int64_t *choose, int32_t *count_out,
                                     void *workspace, size_t workspace_bytes, void *stream) {
  if (frames < 0 || img_h <= 0 || img_w <= 0 || n_points <= 0 || (long long)img_h * img_w > (1ll << 30) || keep_stride < 0) {
    set_error("plr2_choose_from_mask: bad shape F=%d img=%dx%d N=%d", frames, img_h, img_w, n_points);
    return PLR2_E_SHAPE;
  }
  if (frames == 0) return 0;
  if (!mask || !bbox || !choose || !workspace) { set_error("plr2_choose_from_mask: null pointer"); return PLR2_E_NULLPTR; }
  // every crop fits in the image, so img_h*img_w entries per frame always suffice; smaller workspaces are
  // accepted when the caller knows its largest crop
  const size_t per_frame = workspace_bytes / sizeof(int32_t) / (size_t)frames;
  if (per_frame == 0) { set_error("plr2_choose_from_mask: workspace too small"); return PLR2_E_WORKSPACE; }
  const int stride = (int)(per_frame < (size_t)img_h * img_w ? per_frame : (size_t)img_h * img_w);
  if (stride < img_h * img_w) {
    // the kernel cannot report an overflow cheaply: require room for the full image unless the host checked
    // the crops (bbox is device memory here), i.e. insist on the documented size
    set_error("plr2_choose_from_mask: workspace %zu < %zu bytes", workspace_bytes,
              plr2_choose_workspace_bytes(frames, img_h * img_w));
    return PLR2_E_WORKSPACE;
  }
  choose_kernel<<<frames, CH_THREADS, 0, (cudaStream_t)stream>>>(mask, img_h, img_w, bbox, n_points, keep, keep_stride,
                                                               seed, choose, count_out, (int32_t *)workspace, stride);
  return check_launch("plr2_choose_from_mask");
}

extern "C" int plr2_backproject(const void *depth, int depth_is_u16, int img_h, int img_w, const int32_t *bbox,
                                const int64_t *choose, const float *cam, const double *add_t, int frames,
                                int n_points, float *cloud, void *stream) {
  if (frames < 0 || img_h <= 0 || img_w <= 0 || n_points <= 0 || frames > 65535) {
    set_error("plr2_backproject: bad shape F=%d img=%dx%d N=%d", frames, img_h, img_w, n_points);
    return PLR2_E_SHAPE;
  }
  if (frames == 0) return 0;
  if (!depth || !bbox || !choose || !cam || !cloud) { set_error("plr2_backproject: null pointer"); return PLR2_E_NULLPTR; }
  dim3 grid((n_points + 255) / 256, frames);
  cudaStream_t st = (cudaStream_t)stream;
  if (depth_is_u16)
    backproject_kernel<uint16_t><<<grid, 256, 0, st>>>((const uint16_t *)depth, img_h, img_w, bbox, choose, cam, add_t, n_points, cloud);
  else
    backproject_kernel<float><<<grid, 256, 0, st>>>((const float *)depth, img_h, img_w, bbox, choose, cam, add_t, n_points, cloud);
  return check_launch("plr2_backproject");
}

extern "C" int plr2_transform_target(const float *model_points, const double *R, const double *t, const double *add_t,
                                     int frames, int n_mesh, float *target, void *stream) {
  if (frames < 0 || n_mesh <= 0 || frames > 65535) {
    set_error("plr2_transform_target: bad shape F=%d M=%d", frames, n_mesh);
    return PLR2_E_SHAPE;
  }
  if (frames == 0) return 0;
  if (!model_points || !R || !t || !target) { set_error("plr2_transform_target: null pointer"); return PLR2_E_NULLPTR; }
  dim3 grid((n_mesh + 255) / 256, frames);
  transform_target_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(model_points, R, t, add_t, n_mesh, target);
  return check_launch("plr2_transform_target");
}

// ---------------------------------------------------------------------------------------------
// mask = (label == obj) & (depth != 0), bbox = get_bbox(label == obj)  (datasets/ycb/dataset.py:143-145, 251-289)
namespace plr2 {

// snap a side length up to the next entry of border_list = [-1, 40, 80, ..., 680] (dataset.py:247,258-267):
// values strictly between two borders go to the upper one, values ON a border (or beyond the list) stay
__device__ __forceinline__ int snap_border(int b) {
  if (b <= -1 || b >= 680) return b;
  if (b < 40) return 40;                                  // between -1 and 40
  return (b % 40 == 0) ? b : (b / 40 + 1) * 40;
}

template <typename DT>
__global__ void __launch_bounds__(1024)
mask_bbox_kernel(const DT *__restrict__ depth, const uint8_t *__restrict__ label, const int32_t *__restrict__ obj_id,
                 int img_h, int img_w, uint8_t *__restrict__ mask, int32_t *__restrict__ bbox, int32_t *__restrict__ count) {
  __shared__ int s_rmin, s_rmax, s_cmin, s_cmax, s_cnt;
  const int f = blockIdx.x, tid = threadIdx.x;
  if (tid == 0) { s_rmin = img_h; s_rmax = -1; s_cmin = img_w; s_cmax = -1; s_cnt = 0; }
  __syncthreads();
  const int obj = obj_id[f];
  const size_t base = (size_t)f * img_h * img_w;
  const int P = img_h * img_w;
  int rmin = img_h, rmax = -1, cmin = img_w, cmax = -1, cnt = 0;
  for (int p = tid; p < P; p += 1024) {
    const bool lab = (int)label[base + p] == obj;
    const bool m = lab && depth[base + p] != (DT)0;
    mask[base + p] = m ? 1 : 0;
    cnt += m ? 1 : 0;
    if (lab) {
      const int r = p / img_w, c = p - r * img_w;
      rmin = min(rmin, r); rmax = max(rmax, r); cmin = min(cmin, c); cmax = max(cmax, c);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    rmin = min(rmin, __shfl_xor_sync(0xffffffffu, rmin, o)); rmax = max(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
    cmin = min(cmin, __shfl_xor_sync(0xffffffffu, cmin, o)); cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((tid & 31) == 0) {
    atomicMin(&s_rmin, rmin); atomicMax(&s_rmax, rmax); atomicMin(&s_cmin, cmin); atomicMax(&s_cmax, cmax);
    atomicAdd(&s_cnt, cnt);
  }
  __syncthreads();
  if (tid != 0) return;
  if (count) count[f] = s_cnt;
  int r0 = s_rmin, r1 = s_rmax + 1, c0 = s_cmin, c1 = s_cmax + 1;       // rmax += 1; cmax += 1 (:256-257)
  if (s_rmax < 0) { bbox[4 * f] = bbox[4 * f + 1] = bbox[4 * f + 2] = bbox[4 * f + 3] = 0; return; }
  const int r_b = snap_border(r1 - r0), c_b = snap_border(c1 - c0);
  const int cr = (r0 + r1) / 2, cc = (c0 + c1) / 2;                    // int((rmin + rmax) / 2): operands >= 0
  r0 = cr - r_b / 2; r1 = cr + r_b / 2; c0 = cc - c_b / 2; c1 = cc + c_b / 2;
  if (r0 < 0) { r1 += -r0; r0 = 0; }
  if (c0 < 0) { c1 += -c0; c0 = 0; }
  if (r1 > img_h) { r0 -= r1 - img_h; r1 = img_h; }
  if (c1 > img_w) { c0 -= c1 - img_w; c1 = img_w; }
  bbox[4 * f] = r0; bbox[4 * f + 1] = r1; bbox[4 * f + 2] = c0; bbox[4 * f + 3] = c1;
}

}  // namespace plr2

extern "C" int plr2_mask_bbox(const void *depth, int depth_is_u16, const uint8_t *label, const int32_t *obj_id, int frames,
                              int img_h, int img_w, uint8_t *mask, int32_t *bbox, int32_t *count, void *stream) {
  if (frames < 0 || img_h <= 0 || img_w <= 0 || (long long)img_h * img_w > (1ll << 30)) {
    plr2::set_error("plr2_mask_bbox: bad shape F=%d img=%dx%d", frames, img_h, img_w);
    return PLR2_E_SHAPE;
  }
  if (frames == 0) return 0;
  if (!depth || !label || !obj_id || !mask || !bbox) { plr2::set_error("plr2_mask_bbox: null pointer"); return PLR2_E_NULLPTR; }
  cudaStream_t st = (cudaStream_t)stream;
  if (depth_is_u16)
    plr2::mask_bbox_kernel<uint16_t><<<frames, 1024, 0, st>>>((const uint16_t *)depth, label, obj_id, img_h, img_w, mask, bbox, count);
  else
    plr2::mask_bbox_kernel<float><<<frames, 1024, 0, st>>>((const float *)depth, label, obj_id, img_h, img_w, mask, bbox, count);
  return plr2::check_launch("plr2_mask_bbox");
}
