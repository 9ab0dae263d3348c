// Kernel (1) and the small non-GEMM pieces of the PointNet path:
//   * embedding gather at `choose` fused with the first PointNet layer (conv1 3->64, e_conv1 32->64)
//     — lib/network.py:98-102 + :54-55 (PoseNet), :152-153 (PoseRefineNet, emb given)
//   * per-frame average pool (AvgPool1d(num_points), lib/network.py:65) and its backward
//   * object-row selection + sigmoid (lib/network.py:119-130) and its backward
//   * bias-gradient column sums, embedding-gradient scatter, weight pack / unpack
#include "pointnet.cuh"
#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace plr2 {

// ---------------------------------------------------------------------------------------------
// First layer.  One thread per point: gathers its 32 embedding channels (coalesced across the
// ascending `choose` indices as far as the data allows), emits emb in the reference's channel-major
// layout, and writes relu(conv1(x)) | relu(e_conv1(emb)) into columns [0,64) | [64,128) of the
// point-major feature buffer pf (ld 384).
constexpr int FL_THREADS = 128;
constexpr int FL_PITCH = 144;      // bytes per staged row: 128 + 16 pad

template <bool OUT_BF16>
__global__ void __launch_bounds__(FL_THREADS)
first_layer_kernel(const float *__restrict__ c1w, const float *__restrict__ c1b,
                   const float *__restrict__ e1w, const float *__restrict__ e1b,
                   const float *__restrict__ out_img, const int64_t *__restrict__ choose,
                   const float *__restrict__ emb_in, const float *__restrict__ x, int N, int hw,
                   float *__restrict__ pf, __nv_bfloat16 *__restrict__ pfb, long long pf_plane,
                   float *__restrict__ emb_out, float *__restrict__ embT, int skip_e,
                   const int64_t *__restrict__ obj, int num_obj, unsigned int *__restrict__ status, int fp16) {
  __shared__ __align__(16) float sw1[64 * 4];
  __shared__ __align__(16) float swe[64 * 32];
  __shared__ float sb1[64], sbe[64];
  __shared__ __align__(16) uint8_t tile[OUT_BF16 ? FL_THREADS / 32 : 1][2][OUT_BF16 ? 32 * FL_PITCH : 16];
  // weight staging: all loads of a thread in flight at once (16-byte, packed buffer offsets are 128-byte aligned)
  {
    const float4 *g1 = reinterpret_cast<const float4 *>(c1w), *ge = reinterpret_cast<const float4 *>(e1w);
    float4 v[5];
#pragma unroll
    for (int u = 0; u < 5; ++u) {
      const int i = threadIdx.x + u * FL_THREADS;           // 64 + 512 float4 in total
      v[u] = i < 64 ? __ldg(g1 + i) : (i < 64 + 512 ? __ldg(ge + (i - 64)) : make_float4(0.f, 0.f, 0.f, 0.f));
    }
    float bv = threadIdx.x < 64 ? __ldg(c1b + threadIdx.x) : (threadIdx.x < 128 ? __ldg(e1b + threadIdx.x - 64) : 0.f);
#pragma unroll
    for (int u = 0; u < 5; ++u) {
      const int i = threadIdx.x + u * FL_THREADS;
      if (i < 64) reinterpret_cast<float4 *>(sw1)[i] = v[u];
      else if (i < 64 + 512) reinterpret_cast<float4 *>(swe)[i - 64] = v[u];
    }
    if (threadIdx.x < 64) sb1[threadIdx.x] = bv;
    else if (threadIdx.x < 128) sbe[threadIdx.x - 64] = bv;
  }
  __syncthreads();
  const int f = blockIdx.y;
  const int n_raw = blockIdx.x * FL_THREADS + threadIdx.x;
  const bool live = n_raw < N;
  const int n = live ? n_raw : N - 1;               // dead lanes shadow the last point (no stores)
  const size_t row = (size_t)f * N + n;
  float e[32];
  if (skip_e) {                                     // emb branch reused from the previous refine iteration
#pragma unroll
    for (int ch = 0; ch < 32; ++ch) e[ch] = 0.f;
  } else if (choose != nullptr) {
    long long c = choose[row];
    if (c < 0 || c >= hw) {                         // torch.gather raises; a kernel cannot: flag it (plr2_net_input_status), then clamp
      if (status != nullptr && live) atomicOr(status, (unsigned int)PLR2_STATUS_CHOOSE_RANGE);
      c = c < 0 ? 0 : hw - 1;
    }
    const float *src = out_img + (size_t)f * 32 * hw + c;
#pragma unroll
    for (int ch = 0; ch < 32; ++ch) e[ch] = __ldg(src + (size_t)ch * hw);
  } else {
    const float *src = emb_in + (size_t)f * 32 * N + n;
#pragma unroll
    for (int ch = 0; ch < 32; ++ch) e[ch] = __ldg(src + (size_t)ch * N);
  }
  if (emb_out != nullptr && live) {
    float *dst = emb_out + (size_t)f * 32 * N + n;
#pragma unroll
    for (int ch = 0; ch < 32; ++ch) dst[(size_t)ch * N] = e[ch];
  }
  if (embT != nullptr && live) {
    float4 *dst = reinterpret_cast<float4 *>(embT + row * 32);
#pragma unroll
    for (int q = 0; q < 8; ++q) dst[q] = make_float4(e[4 * q], e[4 * q + 1], e[4 * q + 2], e[4 * q + 3]);
  }
  if (status != nullptr && obj != nullptr && n_raw == 0) {      // one thread per frame: class index the heads will select
    const long long o = obj[f];
    if (o < 0 || o >= num_obj) atomicOr(status, (unsigned int)PLR2_STATUS_OBJ_RANGE);
  }
  const float px = x[row * 3], py = x[row * 3 + 1], pz = x[row * 3 + 2];
  // 8 outputs at a time: conv1 (o < 64) then e_conv1 (o >= 64)
  const int o8_end = skip_e ? 8 : 16;
#pragma unroll 1
  for (int o8 = 0; o8 < o8_end; ++o8) {
    float v[8];
    if (o8 < 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int o = o8 * 8 + j;
        const float4 w = *reinterpret_cast<const float4 *>(sw1 + o * 4);
        v[j] = fmaxf(fmaf(w.z, pz, fmaf(w.y, py, fmaf(w.x, px, sb1[o]))), 0.f);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int o = (o8 - 8) * 8 + j;
        float acc = sbe[o];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 w = *reinterpret_cast<const float4 *>(swe + o * 32 + q * 4);
          acc = fmaf(w.x, e[4 * q], acc); acc = fmaf(w.y, e[4 * q + 1], acc);
          acc = fmaf(w.z, e[4 * q + 2], acc); acc = fmaf(w.w, e[4 * q + 3], acc);
        }
        v[j] = fmaxf(acc, 0.f);
      }
    }
    if (OUT_BF16) {
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (fp16) {
          const __half2 h = __floats2half2_rn(v[2 * q], v[2 * q + 1]);
          const float2 hf = __half22float2(h);
          const __half2 l = __floats2half2_rn(v[2 * q] - hf.x, v[2 * q + 1] - hf.y);
          hi[q] = *reinterpret_cast<const uint32_t *>(&h);
          lo[q] = *reinterpret_cast<const uint32_t *>(&l);
        } else {
          const __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * q]), h1 = __float2bfloat16_rn(v[2 * q + 1]);
          const __nv_bfloat16 l0 = __float2bfloat16_rn(v[2 * q] - __bfloat162float(h0));
          const __nv_bfloat16 l1 = __float2bfloat16_rn(v[2 * q + 1] - __bfloat162float(h1));
          hi[q] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          lo[q] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
        }
      }
      // Stage 64 output columns of the warp's 32 points ([32][8 x 16 B], pitch 144 B: conflict-free), then
      // write them out as 128-byte row segments (4 rows per instruction instead of 32 scattered 16-byte pieces).
      const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
      uint8_t *th = tile[wrp][0], *tl = tile[wrp][1];
      *reinterpret_cast<uint4 *>(th + lane * FL_PITCH + (o8 & 7) * 16) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      if (pf_plane != 0) *reinterpret_cast<uint4 *>(tl + lane * FL_PITCH + (o8 & 7) * 16) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      if ((o8 & 7) == 7) {
        __syncwarp();
        const int n_warp0 = blockIdx.x * FL_THREADS + wrp * 32;              // first point of this warp
        const int seg = lane & 7, rsub = lane >> 3;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = 4 * i + rsub;
          if (n_warp0 + r < N) {
            __nv_bfloat16 *d = pfb + ((size_t)f * N + n_warp0 + r) * 384 + (o8 >> 3) * 64 + seg * 8;
            *reinterpret_cast<uint4 *>(d) = *reinterpret_cast<const uint4 *>(th + r * FL_PITCH + seg * 16);
            if (pf_plane != 0)                                               // 0: hi-only consumers
              *reinterpret_cast<uint4 *>(d + pf_plane) = *reinterpret_cast<const uint4 *>(tl + r * FL_PITCH + seg * 16);
          }
        }
        __syncwarp();
      }
    } else if (live) {
      float4 *d = reinterpret_cast<float4 *>(pf + row * 384 + o8 * 8);
      d[0] = make_float4(v[0], v[1], v[2], v[3]);
      d[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
  }
}

int first_layer(const float *packed, const NetLayout &L, const float *out_img, const int64_t *choose,
                const float *emb_in, const float *x, int frames, int N, int hw, float *pf, float *emb_out,
                float *embT, const int64_t *obj, unsigned int *status, cudaStream_t st) {
  dim3 grid((N + FL_THREADS - 1) / FL_THREADS, frames);
  first_layer_kernel<false><<<grid, FL_THREADS, 0, st>>>(packed + L.c1w, packed + L.c1b, packed + L.e1w, packed + L.e1b,
                                                         out_img, choose, emb_in, x, N, hw, pf, nullptr, 0, emb_out, embT, 0,
                                                         obj, L.O, status, 0);
  return check_launch("first_layer");
}

// same, writing the split-bf16 planes consumed by the tensor-core layers
int first_layer_bf16(const float *packed, const NetLayout &L, const float *out_img, const int64_t *choose,
                     const float *emb_in, const float *x, int frames, int N, int hw, __nv_bfloat16 *pfb,
                     long long pf_plane, float *emb_out, int skip_e, const int64_t *obj, unsigned int *status,
                     cudaStream_t st) {
  dim3 grid((N + FL_THREADS - 1) / FL_THREADS, frames);
  first_layer_kernel<true><<<grid, FL_THREADS, 0, st>>>(packed + L.c1w, packed + L.c1b, packed + L.e1w, packed + L.e1b,
                                                        out_img, choose, emb_in, x, N, hw, nullptr, pfb, pf_plane, emb_out,
                                                        nullptr, skip_e, obj, L.O, status, infer_fp16() ? 1 : 0);
  return check_launch("first_layer_bf16");
}

// d out_img[f][c][choose[f][n]] += demb[row][c]  (duplicates in choose accumulate)
__global__ void __launch_bounds__(256)
scatter_emb_grad_kernel(const float *__restrict__ demb, const int64_t *__restrict__ choose, int N, int hw,
                        size_t rows, float *__restrict__ g_out_img) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * 32) return;
  const size_t row = i >> 5;
  const int c = (int)(i & 31);
  const int f = (int)(row / N);
  long long p = choose[row];
  p = p < 0 ? 0 : (p >= hw ? hw - 1 : p);
  atomicAdd(g_out_img + ((size_t)f * 32 + c) * hw + p, demb[i]);
}

int scatter_emb_grad(const float *demb, const int64_t *choose, int frames, int N, int hw, float *g_out_img,
                     cudaStream_t st) {
  const size_t rows = (size_t)frames * N;
  scatter_emb_grad_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, st>>>(demb, choose, N, hw, rows, g_out_img);
  return check_launch("scatter_emb_grad");
}

// ---------------------------------------------------------------------------------------------
// out[seg][c] = scale * sum_{r in segment} X[r][c]   (segments of rows_per_seg consecutive rows).
// Block: 32 columns x 8 row lanes; fixed summation order (deterministic).
__global__ void __launch_bounds__(256)
segment_colsum_kernel(const float *__restrict__ X, long long ld, int rows_per_seg, size_t total_rows, int ncols, float scale,
                      float *__restrict__ out, long long ldo) {
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rl = threadIdx.x >> 5;
  const size_t r0 = (size_t)blockIdx.y * rows_per_seg;
  const int nr = (int)min((size_t)rows_per_seg, total_rows > r0 ? total_rows - r0 : (size_t)0);   // last segment may be short
  float s = 0.f;
  if (c < ncols) {
    // 4 independent partial sums per thread (rows rl, rl+8, rl+16, rl+24, ...): loads in flight instead of a
    // serial add chain; combined in a fixed order
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int r = rl;
    for (; r + 24 < nr; r += 32) {
      s0 += X[(r0 + r) * ld + c]; s1 += X[(r0 + r + 8) * ld + c];
      s2 += X[(r0 + r + 16) * ld + c]; s3 += X[(r0 + r + 24) * ld + c];
    }
    for (; r < nr; r += 8) s0 += X[(r0 + r) * ld + c];
    s = (s0 + s1) + (s2 + s3);
  }
  red[rl][threadIdx.x & 31] = s;
  __syncthreads();
  if (rl == 0 && c < ncols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    out[(size_t)blockIdx.y * ldo + c] = t * scale;
  }
}

// One segment of a few hundred rows (the per-tile partials of a bias gradient): 32 row lanes per block instead
// of 8, so that a thread's chain is rows/32 loads long.  Fixed summation order.
__global__ void __launch_bounds__(1024)
colsum_tall_kernel(const float *__restrict__ X, long long ld, int rows, int ncols, float scale, float *__restrict__ out) {
  __shared__ float red[32][33];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rl = threadIdx.x >> 5;
  float s0 = 0.f, s1 = 0.f;
  if (c < ncols) {
    int r = rl;
    for (; r + 32 < rows; r += 64) { s0 += X[(size_t)r * ld + c]; s1 += X[(size_t)(r + 32) * ld + c]; }
    if (r < rows) s0 += X[(size_t)r * ld + c];
  }
  red[rl][threadIdx.x & 31] = s0 + s1;
  __syncthreads();
  if (rl == 0 && c < ncols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) t += red[i][threadIdx.x];
    out[c] = t * scale;
  }
}

int colsum_tall(const float *X, long long ld, int rows, int ncols, float *out, cudaStream_t st) {
  if (rows <= 0 || ncols <= 0) return 0;
  colsum_tall_kernel<<<(ncols + 31) / 32, 1024, 0, st>>>(X, ld, rows, ncols, 1.0f, out);
  return check_launch("colsum_tall");
}

int segment_colsum(const float *X, long long ld, int nseg, int rows_per_seg, int ncols, float scale, float *out,
                   long long ldo, cudaStream_t st) {
  if (nseg <= 0 || ncols <= 0) return 0;

  dim3 grid((ncols + 31) / 32, nseg);
  segment_colsum_kernel<<<grid, 256, 0, st>>>(X, ld, rows_per_seg, (size_t)nseg * rows_per_seg, ncols, scale, out, ldo);
  return check_launch("segment_colsum");
}

// same with a short last segment: segment s covers rows [s*rows_per_seg, min((s+1)*rows_per_seg, total_rows))
int segment_colsum_ragged(const float *X, long long ld, int nseg, int rows_per_seg, size_t total_rows, int ncols, float *out,
                          long long ldo, cudaStream_t st) {
  if (nseg <= 0 || ncols <= 0) return 0;
  dim3 grid((ncols + 31) / 32, nseg);
  segment_colsum_kernel<<<grid, 256, 0, st>>>(X, ld, rows_per_seg, total_rows, ncols, 1.0f, out, ldo);
  return check_launch("segment_colsum_ragged");
}

// dX[r][c] = (X[r][c] > 0) * dseg[r / rows_per_seg][c] * scale     (in place over X allowed)
__global__ void __launch_bounds__(256)
segment_bcast_mask_kernel(const float *__restrict__ dseg, const float *X, int rows_per_seg, int ncols,
                          size_t total, float scale, float *dX) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= total) return;
  const size_t r = i / ncols;
  const int c = (int)(i % ncols);
  const float4 x = *reinterpret_cast<const float4 *>(X + i);
  const float4 g = *reinterpret_cast<const float4 *>(dseg + (r / rows_per_seg) * ncols + c);
  float4 o;
  o.x = x.x > 0.f ? g.x * scale : 0.f; o.y = x.y > 0.f ? g.y * scale : 0.f;
  o.z = x.z > 0.f ? g.z * scale : 0.f; o.w = x.w > 0.f ? g.w * scale : 0.f;
  *reinterpret_cast<float4 *>(dX + i) = o;
}

int segment_bcast_mask(const float *dseg, const float *X, size_t rows, int rows_per_seg, int ncols, float scale,
                       float *dX, cudaStream_t st) {
  const size_t total = rows * ncols;   // ncols is a multiple of 4 (1024)
  segment_bcast_mask_kernel<<<(unsigned)((total / 4 + 255) / 256), 256, 0, st>>>(dseg, X, rows_per_seg, ncols,
                                                                               total, scale, dX);
  return check_launch("segment_bcast_mask");
}

// ---------------------------------------------------------------------------------------------
// Object selection: raw [rows][ldo] holds every class's outputs; pick the class of the row's frame.
// (lib/network.py:119-130 / :199-204).  pred_c gets the sigmoid.
__global__ void __launch_bounds__(256)
select_outputs_kernel(const float *__restrict__ raw, int ldo, int oc_r, int oc_t, int oc_c,
                      const int64_t *__restrict__ obj, int num_obj, int rows_per_frame, size_t rows,
                      float *__restrict__ pred_r, float *__restrict__ pred_t, float *__restrict__ pred_c) {
  const size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  long long o = obj[r / rows_per_frame];
  o = o < 0 ? 0 : (o >= num_obj ? num_obj - 1 : o);
  const float *src = raw + r * ldo;
#pragma unroll
  for (int i = 0; i < 4; ++i) pred_r[r * 4 + i] = src[oc_r + o * 4 + i];
#pragma unroll
  for (int i = 0; i < 3; ++i) pred_t[r * 3 + i] = src[oc_t + o * 3 + i];
  if (pred_c != nullptr) pred_c[r] = 1.0f / (1.0f + expf(-src[oc_c + o]));
}

int select_outputs(const float *raw, const NetLayout &L, const int64_t *obj, int rows_per_frame, size_t rows,
                   float *pred_r, float *pred_t, float *pred_c, cudaStream_t st) {
  select_outputs_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(raw, L.ldo, L.oc[0], L.oc[1], L.oc[2], obj,
                                                                       L.O, rows_per_frame, rows, pred_r, pred_t, pred_c);
  return check_launch("select_outputs");
}

// Backward of the selection, written IN PLACE over the raw buffer: zero everywhere except the
// selected class's columns; the confidence column gets g * s(1-s) recomputed from its logit.
__global__ void __launch_bounds__(256)
select_outputs_bwd_kernel(float *raw, int ldo, int oc_r, int oc_t, int oc_c, const int64_t *__restrict__ obj,
                          int num_obj, int rows_per_frame, size_t rows, const float *__restrict__ g_r,
                          const float *__restrict__ g_t, const float *__restrict__ g_c) {
  const size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  long long o = obj[r / rows_per_frame];
  o = o < 0 ? 0 : (o >= num_obj ? num_obj - 1 : o);
  float *dst = raw + r * ldo;
  float gc = 0.f;
  if (g_c != nullptr) {
    const float s = 1.0f / (1.0f + expf(-dst[oc_c + o]));
    gc = g_c[r] * s * (1.0f - s);
  }
  for (int i = 0; i < ldo; ++i) dst[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) dst[oc_r + o * 4 + i] = g_r[r * 4 + i];
#pragma unroll
  for (int i = 0; i < 3; ++i) dst[oc_t + o * 3 + i] = g_t[r * 3 + i];
  if (g_c != nullptr) dst[oc_c + o] = gc;
}

int select_outputs_bwd(float *raw, const NetLayout &L, const int64_t *obj, int rows_per_frame, size_t rows,
                       const float *g_r, const float *g_t, const float *g_c, cudaStream_t st) {
  select_outputs_bwd_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(raw, L.ldo, L.oc[0], L.oc[1], L.oc[2], obj,
                                                                           L.O, rows_per_frame, rows, g_r, g_t, g_c);
  return check_launch("select_outputs_bwd");
}

// ---------------------------------------------------------------------------------------------
// Weight pack / unpack: 2-D block copies described by a small table passed by value.
struct CopyEntry {
  const float *src; float *dst;
  long long src_ld, dst_ld;
  int rows, cols, round;
};
constexpr int MAX_ENTRIES = 48;
struct CopyTable {
  CopyEntry e[MAX_ENTRIES];
  int n;
  const float *base;            // packed FP32 region (for the element offset of a destination)
  __nv_bfloat16 *planes;        // split hi/lo planes of the same layout (bf16 or fp16 bits), or nullptr
  long long plane_stride;
  int fp16;
  int accumulate;               // dst += src (gradient unpack onto existing .grad tensors)
};

__global__ void __launch_bounds__(256) copy_blocks_kernel(const __grid_constant__ CopyTable T) {
  const CopyEntry &E = T.e[blockIdx.y];
  const size_t total = (size_t)E.rows * E.cols;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / E.cols, c = i % E.cols;
    float v = E.src[r * E.src_ld + c];
    if (T.accumulate) v += E.dst[r * E.dst_ld + c];
    E.dst[r * E.dst_ld + c] = v;
    if (E.round && T.planes != nullptr) {     // split-bf16 copy for the tensor-core layers
      const long long off = (E.dst - T.base) + (long long)(r * E.dst_ld + c);
      if (T.fp16) {
        const __half h = __float2half_rn(v);
        __half *pl = reinterpret_cast<__half *>(T.planes);
        pl[off] = h;
        pl[T.plane_stride + off] = __float2half_rn(v - __half2float(h));
      } else {
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        T.planes[off] = h;
        T.planes[T.plane_stride + off] = __float2bfloat16_rn(v - __bfloat162float(h));
      }
    }
  }
}

static void add_entry(CopyTable &T, const float *src, long long src_ld, float *dst, long long dst_ld, int rows,
                      int cols, int round) {
  if (rows <= 0 || cols <= 0) return;
  CopyEntry &E = T.e[T.n++];
  E.src = src; E.dst = dst; E.src_ld = src_ld; E.dst_ld = dst_ld; E.rows = rows; E.cols = cols; E.round = round;
}

// Builds the table mapping state-dict tensors <-> packed layout.  `to_packed`: params are the
// source; otherwise the packed buffer is the source (gradient unpack).
static int build_table(const NetLayout &L, const plr2_param_table *P, float *packed, bool to_packed, int round,
                       CopyTable &T) {
  T.n = 0;
  const bool pn = L.net == PLR2_NET_POSENET;
  const int nl = pn ? PLR2_POSENET_LAYERS : PLR2_REFINER_LAYERS;
  for (int i = 0; i < nl; ++i)
    if (!P->w[i] || !P->b[i]) { set_error("net pack: parameter %d is NULL", i); return PLR2_E_NULLPTR; }
  auto blk = [&](float *param, long long pld, size_t off, long long kld, int rows, int cols, int rnd) {
    if (to_packed) add_entry(T, param, pld, packed + off, kld, rows, cols, rnd);
    else add_entry(T, packed + off, kld, param, pld, rows, cols, 0);
  };
  // feature layers 0..5: conv1, e_conv1, conv2, e_conv2, conv5, conv6
  blk(P->w[0], 3, L.c1w, 4, 64, 3, 0);           blk(P->b[0], 64, L.c1b, 64, 1, 64, 0);
  blk(P->w[1], 32, L.e1w, 32, 64, 32, 0);        blk(P->b[1], 64, L.e1b, 64, 1, 64, 0);
  blk(P->w[2], 64, L.c2w, 64, 128, 64, round);   blk(P->b[2], 128, L.c2b, 128, 1, 128, 0);
  blk(P->w[3], 64, L.e2w, 64, 128, 64, round);   blk(P->b[3], 128, L.e2b, 128, 1, 128, 0);
  blk(P->w[4], L.c5k, L.c5w, L.c5k, 512, L.c5k, round);  blk(P->b[4], 512, L.c5b, 512, 1, 512, 0);
  blk(P->w[5], 512, L.c6w, 512, 1024, 512, round);       blk(P->b[5], 1024, L.c6b, 1024, 1, 1024, 0);
  int li = 6;
  for (int h = 0; h < L.heads; ++h, ++li) {      // conv1_{r,t,c}
    if (pn) {
      blk(P->w[li], 1408, L.h1pf + (size_t)h * 640 * 384, 384, 640, 384, round);
      blk(P->w[li] + 384, 1408, L.h1ap + (size_t)h * 640 * 1024, 1024, 640, 1024, round);
    } else {
      blk(P->w[li], 1024, L.h1ap + (size_t)h * 512 * 1024, 1024, 512, 1024, round);
    }
    blk(P->b[li], L.h1n, L.h1b + (size_t)h * L.h1n, L.h1n, 1, L.h1n, 0);
  }
  for (int h = 0; h < L.heads; ++h, ++li) {      // conv2_*
    blk(P->w[li], L.h2k, L.h2w + (size_t)h * L.h2n * L.h2k, L.h2k, L.h2n, L.h2k, pn ? round : 0);
    blk(P->b[li], L.h2n, L.h2b + (size_t)h * L.h2n, L.h2n, 1, L.h2n, 0);
  }
  if (pn)
    for (int h = 0; h < 3; ++h, ++li) {          // conv3_* (PoseNet only)
      blk(P->w[li], 256, L.h3w + (size_t)h * 128 * 256, 256, 128, 256, round);
      blk(P->b[li], 128, L.h3b + (size_t)h * 128, 128, 1, 128, 0);
    }
  size_t row = 0;
  for (int h = 0; h < L.heads; ++h, ++li) {      // last layer (conv4_* / refiner conv3_*)
    blk(P->w[li], 128, L.h4w + row * 128, 128, L.on[h], 128, pn ? round : 0);
    blk(P->b[li], L.on[h], L.h4b + row, L.on[h], 1, L.on[h], 0);
    row += L.on[h];
  }
  return 0;
}

int net_copy_params(const NetLayout &L, const plr2_param_table *P, float *packed, bool to_packed, int round,
                    cudaStream_t st, int accumulate) {
  CopyTable T;
  T.accumulate = accumulate;
  int rc = build_table(L, P, packed, to_packed, round, T);
  if (rc) return rc;
  T.base = packed;
  T.planes = (to_packed && round) ? reinterpret_cast<__nv_bfloat16 *>(packed + L.total) : nullptr;
  T.plane_stride = (long long)L.total;
  T.fp16 = infer_fp16() ? 1 : 0;   // the packed planes feed the inference path only (training packs its operands per contraction)
  if (to_packed) {
    // conv1 weight is padded from 3 to 4 columns: clear the pad lane once
    cudaError_t e = cudaMemsetAsync(packed + L.c1w, 0, 64 * 4 * sizeof(float), st);
    if (e != cudaSuccess) { set_error("net pack memset: %s", cudaGetErrorString(e)); return (int)e; }
  }
  dim3 grid(64, T.n);
  copy_blocks_kernel<<<grid, 256, 0, st>>>(T);
  return check_launch("net_copy_params");
}

}  // namespace plr2
