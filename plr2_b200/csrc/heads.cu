// Small FP32 SIMT kernels around the tensor-core layers of kernel (2):
//   * frame_linear: Y[M,N] = act(X[M,K] W[N,K]^T + b) for M = a few frames (the per-frame bias of
//     conv1_{r,t,c}, lib/network.py:107-109 after the hoist of SURVEY.md §7; the Linear layers of
//     PoseRefineNet, lib/network.py:193-197).  The weights are streamed once per 64 frames.
//   * head_out: the LAST layer of each head evaluated only for the object class of the row's frame
//     (conv4_{r,t,c} + index_select + sigmoid, lib/network.py:117-130; refiner conv3_{r,t} + select,
//     :199-204): 8 (7) dot products of length 128 per row instead of 8*num_obj.
#include "pointnet.cuh"

namespace plr2 {

// ---------------------------------------------------------------------------------------------
constexpr int FLN_ROWS = 64;      // rows (frames) per CTA: 2 per lane
constexpr int FLN_COLS = 16;      // output columns per CTA: 4 per warp of a K group
constexpr int FLN_KC = 32;        // K chunk staged in shared memory
constexpr int FLN_KG = 4;         // K groups per CTA (4 warps each), partial sums combined at the end
constexpr int FLN_GT = 128;       // threads per K group
constexpr int FLN_THREADS = FLN_KG * FLN_GT;

struct FrameLinearArgs {
  const float *X; long long ldx;
  const float *W; long long ldw;
  const float *bias;
  float *Y; long long ldy;
  int M, N, K, relu;
  long long zX, zW, zB, zY;        // per-batch (blockIdx.z) offsets in elements
};

__global__ void __launch_bounds__(FLN_THREADS)
frame_linear_kernel(const FrameLinearArgs g) {
  __shared__ __align__(16) float Xs[FLN_KG][FLN_KC][FLN_ROWS + 1];   // [group][k][row]
  __shared__ __align__(16) float Ws[FLN_KG][FLN_KC][FLN_COLS];       // [group][k][col]
  const int grp = threadIdx.x / FLN_GT, tid = threadIdx.x - grp * FLN_GT;
  const int lane = tid & 31, warp = tid >> 5;
  const int n0 = blockIdx.x * FLN_COLS, m0 = blockIdx.y * FLN_ROWS, z = blockIdx.z;
  const float *X = g.X + z * g.zX, *W = g.W + z * g.zW;
  float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
  // this group's K range: whole chunks, split evenly
  const int nchunks = (g.K + FLN_KC - 1) / FLN_KC;
  const int c_beg = nchunks * grp / FLN_KG, c_end = nchunks * (grp + 1) / FLN_KG;

  // global -> registers for one K chunk: X tile 64 rows x 32 k = 512 float4 (4 per thread),
  // W tile 16 cols x 32 k = 128 float4 (1 per thread); rows are K-contiguous in both.
  constexpr int XV = FLN_ROWS * FLN_KC / 4 / FLN_GT, WV = FLN_COLS * FLN_KC / 4 / FLN_GT, KQ = FLN_KC / 4;
  float4 xr[XV], wr[WV];
  const bool vec = ((g.ldx | g.ldw | g.zX | g.zW) & 3) == 0 && (g.K & 3) == 0 &&
                   ((reinterpret_cast<uintptr_t>(g.X) | reinterpret_cast<uintptr_t>(g.W)) & 15) == 0;
  auto load_chunk = [&](int k0) {
#pragma unroll
    for (int i = 0; i < XV; ++i) {
      const int e = tid + i * FLN_GT, r = e / KQ, kq = (e % KQ) * 4;
      const int m = m0 + r, k = k0 + kq;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < g.M) {
        const float *p = X + (long long)m * g.ldx + k;
        if (vec && k + 3 < g.K) v = *reinterpret_cast<const float4 *>(p);
        else {
          if (k < g.K) v.x = p[0];
          if (k + 1 < g.K) v.y = p[1];
          if (k + 2 < g.K) v.z = p[2];
          if (k + 3 < g.K) v.w = p[3];
        }
      }
      xr[i] = v;
    }
#pragma unroll
    for (int i = 0; i < WV; ++i) {
      const int e = tid + i * FLN_GT, c = e / KQ, kq = (e % KQ) * 4;
      const int n = n0 + c, k = k0 + kq;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n < g.N) {
        const float *p = W + (long long)n * g.ldw + k;
        if (vec && k + 3 < g.K) v = __ldg(reinterpret_cast<const float4 *>(p));
        else {
          if (k < g.K) v.x = p[0];
          if (k + 1 < g.K) v.y = p[1];
          if (k + 2 < g.K) v.z = p[2];
          if (k + 3 < g.K) v.w = p[3];
        }
      }
      wr[i] = v;
    }
  };
  auto store_chunk = [&]() {
#pragma unroll
    for (int i = 0; i < XV; ++i) {
      const int e = tid + i * FLN_GT, r = e / KQ, kq = (e % KQ) * 4;
      Xs[grp][kq][r] = xr[i].x; Xs[grp][kq + 1][r] = xr[i].y; Xs[grp][kq + 2][r] = xr[i].z; Xs[grp][kq + 3][r] = xr[i].w;
    }
#pragma unroll
    for (int i = 0; i < WV; ++i) {
      const int e = tid + i * FLN_GT, c = e / KQ, kq = (e % KQ) * 4;
      Ws[grp][kq][c] = wr[i].x; Ws[grp][kq + 1][c] = wr[i].y; Ws[grp][kq + 2][c] = wr[i].z; Ws[grp][kq + 3][c] = wr[i].w;
    }
  };
  auto group_sync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(FLN_GT) : "memory"); };

  if (c_beg < c_end) load_chunk(c_beg * FLN_KC);
  for (int c = c_beg; c < c_end; ++c) {
    store_chunk();
    group_sync();
    if (c + 1 < c_end) load_chunk((c + 1) * FLN_KC);       // in flight while this chunk is consumed
#pragma unroll 8
    for (int k = 0; k < FLN_KC; ++k) {
      const float x0 = Xs[grp][k][lane], x1 = Xs[grp][k][lane + 32];
      const float4 w = *reinterpret_cast<const float4 *>(&Ws[grp][k][warp * 4]);
      acc0[0] = fmaf(x0, w.x, acc0[0]); acc0[1] = fmaf(x0, w.y, acc0[1]);
      acc0[2] = fmaf(x0, w.z, acc0[2]); acc0[3] = fmaf(x0, w.w, acc0[3]);
      acc1[0] = fmaf(x1, w.x, acc1[0]); acc1[1] = fmaf(x1, w.y, acc1[1]);
      acc1[2] = fmaf(x1, w.z, acc1[2]); acc1[3] = fmaf(x1, w.w, acc1[3]);
    }
    group_sync();
  }
  // combine the K groups through shared memory in fixed order (group 0 + 1 + 2 + 3)
  __syncthreads();
  float *part = &Xs[0][0][0];                               // (FLN_KG - 1) x 8 x 128 floats <= sizeof(Xs)
  if (grp > 0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      part[((grp - 1) * 8 + j * 2) * FLN_GT + tid] = acc0[j];
      part[((grp - 1) * 8 + j * 2 + 1) * FLN_GT + tid] = acc1[j];
    }
  }
  __syncthreads();
  if (grp != 0) return;
#pragma unroll
  for (int gq = 1; gq < FLN_KG; ++gq)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      acc0[j] += part[((gq - 1) * 8 + j * 2) * FLN_GT + tid];
      acc1[j] += part[((gq - 1) * 8 + j * 2 + 1) * FLN_GT + tid];
    }
  const float *bias = g.bias ? g.bias + z * g.zB : nullptr;
  float *Y = g.Y + z * g.zY;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int m = m0 + lane + 32 * h;
    if (m >= g.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + warp * 4 + j;
      if (n >= g.N) continue;
      float v = (h ? acc1[j] : acc0[j]) + (bias ? bias[n] : 0.f);
      if (g.relu) v = fmaxf(v, 0.f);
      Y[(long long)m * g.ldy + n] = v;
    }
  }
}

int frame_linear(const float *X, long long ldx, const float *W, long long ldw, const float *bias, float *Y,
                 long long ldy, int M, int N, int K, int relu, int batch, long long zX, long long zW, long long zB,
                 long long zY, cudaStream_t st) {
  if (M <= 0 || N <= 0 || batch <= 0) return 0;
  if (K <= 0) { set_error("frame_linear: K must be positive"); return PLR2_E_SHAPE; }
  FrameLinearArgs g{X, ldx, W, ldw, bias, Y, ldy, M, N, K, relu, zX, zW, zB, zY};
  dim3 grid((N + FLN_COLS - 1) / FLN_COLS, (M + FLN_ROWS - 1) / FLN_ROWS, batch);
  frame_linear_kernel<<<grid, FLN_THREADS, 0, st>>>(g);
  return check_launch("frame_linear");
}

// ---------------------------------------------------------------------------------------------
// head_out: H [rows][ldh] holds the 128-wide inputs of the last layer of each head side by side
// (r | t | c).  W4 [(4+3+1)*O][128] stacked r rows, t rows, c rows (net_layout.cuh), b4 likewise.
// One warp per row; lane owns k = 4*lane .. 4*lane+3 of every head.  The 8 partial dot products
// are reduced with a halving butterfly (4+2+1 exchanges, then 2 plain steps).
constexpr int HO_THREADS = 256;
constexpr int HO_ROWS = 64;       // rows per CTA (8 per warp), all of one frame

template <bool HAS_C>
__global__ void __launch_bounds__(HO_THREADS)
head_out_kernel(const float *__restrict__ H, long long ldh, const float *__restrict__ W4,
                const float *__restrict__ b4, const int64_t *__restrict__ obj, int num_obj, int rows_per_frame,
                float *__restrict__ pred_r, float *__restrict__ pred_t, float *__restrict__ pred_c) {
  const int f = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long o = obj[f];
  o = o < 0 ? 0 : (o >= num_obj ? num_obj - 1 : o);
  // rows of W4 for this frame's class: r: 4o..4o+3 | t: 4O+3o.. | c: 7O+o
  int wrow[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) wrow[j] = (int)(4 * o + j);
#pragma unroll
  for (int j = 0; j < 3; ++j) wrow[4 + j] = (int)(4 * num_obj + 3 * o + j);
  wrow[7] = HAS_C ? (int)(7 * num_obj + o) : 0;
  float4 w[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) w[j] = __ldg(reinterpret_cast<const float4 *>(W4 + (size_t)wrow[j] * 128) + lane);
  // the lane that ends up holding output j after the butterfly: lane = 4*j' pattern below
  const int jout = lane >> 2;                       // output index held by this lane at the end
  const float bj = ((lane & 3) == 0 && (HAS_C || jout < 7)) ? __ldg(b4 + wrow[jout]) : 0.f;

  const int r_beg = blockIdx.x * HO_ROWS;
  for (int rr = warp; rr < HO_ROWS; rr += HO_THREADS / 32) {
    const int n = r_beg + rr;
    if (n >= rows_per_frame) break;
    const size_t row = (size_t)f * rows_per_frame + n;
    const float *h = H + row * ldh;
    const float4 hr = *(reinterpret_cast<const float4 *>(h) + lane);
    const float4 ht = *(reinterpret_cast<const float4 *>(h + 128) + lane);
    float4 hc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (HAS_C) hc = *(reinterpret_cast<const float4 *>(h + 256) + lane);
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 x = j < 4 ? hr : (j < 7 ? ht : hc);
      v[j] = fmaf(x.w, w[j].w, fmaf(x.z, w[j].z, fmaf(x.y, w[j].y, x.x * w[j].x)));
    }
    // halving butterfly: after step s the lane keeps the half of the values selected by its bit
    {
      const bool up = lane & 16;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float send = up ? v[j] : v[j + 4];
        const float keep = up ? v[j + 4] : v[j];
        v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
      }
    }
    {
      const bool up = lane & 8;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float send = up ? v[j] : v[j + 2];
        const float keep = up ? v[j + 2] : v[j];
        v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
      }
    }
    {
      const bool up = lane & 4;
      const float send = up ? v[0] : v[1];
      const float keep = up ? v[1] : v[0];
      v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    // lane bits (16,8,4) = (b2,b1,b0) of the output index j = 4*b2 + 2*b1 + b0  == lane >> 2
    if ((lane & 3) == 0) {
      const float y = v[0] + bj;
      if (jout < 4) pred_r[row * 4 + jout] = y;
      else if (jout < 7) pred_t[row * 3 + (jout - 4)] = y;
      else if (HAS_C) pred_c[row] = 1.0f / (1.0f + expf(-y));
    }
  }
}

int head_out(const float *H, long long ldh, const float *W4, const float *b4, const int64_t *obj, int num_obj,
             int frames, int rows_per_frame, float *pred_r, float *pred_t, float *pred_c, cudaStream_t st) {
  if (frames <= 0 || rows_per_frame <= 0) return 0;
  if ((ldh & 3) || (reinterpret_cast<uintptr_t>(H) & 15) || (reinterpret_cast<uintptr_t>(W4) & 15)) {
    set_error("head_out: operands must be 16-byte aligned");
    return PLR2_E_ALIGN;
  }
  dim3 grid((rows_per_frame + HO_ROWS - 1) / HO_ROWS, frames);
  if (pred_c != nullptr)
    head_out_kernel<true><<<grid, HO_THREADS, 0, st>>>(H, ldh, W4, b4, obj, num_obj, rows_per_frame, pred_r, pred_t, pred_c);
  else
    head_out_kernel<false><<<grid, HO_THREADS, 0, st>>>(H, ldh, W4, b4, obj, num_obj, rows_per_frame, pred_r, pred_t, pred_c);
  return check_launch("head_out");
}

}  // namespace plr2
