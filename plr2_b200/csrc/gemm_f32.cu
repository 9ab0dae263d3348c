// FP32 SIMT GEMM with fused epilogues: the exact-precision path for every contraction of the
// PointNet / head MLPs (lib/network.py:53-68,107-121,151-168,193-200) and of their backward
// (dX = dY.W, dW = dY^T.X).  Inference runs the large layers on the tcgen05 split-bf16 kernel
// (gemm_bf16x2.cu); this kernel serves training (forward and backward), the FP32 precision mode and is the
// in-GPU reference for the tensor-core path.
//
//   C(m,n) (+)= epilogue( sum_k A(m,k) * B(k,n) )        A(m,k)=A[m*sam+k*sak], B(k,n)=B[k*sbk+n*sbn]
//   epilogue: + bias[n] + frame_bias[m / rows_per_frame][n], ReLU, * (relu_mask(m,n) > 0)
//
// 128x128x8 block tile, 256 threads, 8x8 register micro-tile held as packed FP32x2 pairs (FFMA2), register-staged
// double-buffered shared memory.
#include "gemm.cuh"

namespace plr2 {

constexpr int GB_M = 128, GB_N = 128, GB_K = 8, G_THREADS = 256;

enum { LOAD_KVEC = 0, LOAD_MVEC = 1, LOAD_SCALAR = 2 };

// Loads this thread's 4 elements of a (GB_K x 128) operand tile (smem layout [k][mn]).  `mn0` is the first
// row/column of the tile, s_mn / s_k the strides of that index in global memory.  The thread's base pointer and
// row validity are computed once; each k step only advances the pointer (no 64-bit multiplies in the main loop).
template <int MODE>
struct TileLoader {
  const float *p;        // address of this thread's first element for the current k tile
  long long step;        // pointer advance per k tile
  long long s_mn, s_k;
  int kofs;              // this thread's k offset inside a tile
  int mn, MN;            // first mn index of this thread, bound
  bool full;             // all 4 mn indices (MVEC/SCALAR) or the row (KVEC) are in range

  __device__ __forceinline__ void init(const float *P, long long s_mn_, long long s_k_, int mn0, int MN_, int kbeg, int tid) {
    s_mn = s_mn_; s_k = s_k_; MN = MN_;
    if (MODE == LOAD_KVEC) {          // k contiguous: one float4 along k for row tid/2
      mn = mn0 + (tid >> 1); kofs = (tid & 1) * 4;
      full = mn < MN;
      p = P + (long long)(full ? mn : 0) * s_mn + kbeg + kofs;
      step = GB_K;
    } else {                          // MVEC: mn contiguous, one float4 along mn for k = tid/32; SCALAR: any strides
      mn = mn0 + (tid & 31) * 4; kofs = tid >> 5;
      full = mn + 3 < MN;
      p = P + (long long)(kbeg + kofs) * s_k + (long long)(mn < MN ? mn : 0) * s_mn;
      step = (long long)GB_K * s_k;
    }
  }
  // k0: first k of the tile being loaded
  __device__ __forceinline__ void load(int k0, int kend, float (&r)[4]) {
    const int k = k0 + kofs;
    if (MODE == LOAD_KVEC) {
      if (full && k + 3 < kend) {
        const float4 v = *reinterpret_cast<const float4 *>(p);
        r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = v.w;
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) r[i] = (full && k + i < kend) ? p[i] : 0.f;
      }
    } else if (MODE == LOAD_MVEC) {
      if (full && k < kend) {
        const float4 v = *reinterpret_cast<const float4 *>(p);
        r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = v.w;
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) r[i] = (k < kend && mn + i < MN) ? p[i] : 0.f;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) r[i] = (k < kend && mn + i < MN) ? p[(long long)i * s_mn] : 0.f;
    }
    p += step;
  }
};

template <int MODE>
__device__ __forceinline__ void store_tile_smem(float (*S)[GB_M + 4], const float (&r)[4], int tid) {
  if (MODE == LOAD_KVEC) {
    const int row = tid >> 1, kq = (tid & 1) * 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) S[kq + i][row] = r[i];
  } else {
    const int kk = tid >> 5, c = (tid & 31) * 4;
    *reinterpret_cast<float4 *>(&S[kk][c]) = make_float4(r[0], r[1], r[2], r[3]);
  }
}

template <int AMODE, int BMODE>
__global__ void __launch_bounds__(G_THREADS, 2)
gemm_f32_kernel(GemmArgs g) {
  __shared__ __align__(16) float As[2][GB_K][GB_M + 4];
  __shared__ __align__(16) float Bs[2][GB_K][GB_N + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * GB_M, n0 = blockIdx.x * GB_N;
  const int zb = blockIdx.z / g.splits, ks = blockIdx.z % g.splits;
  g.A += zb * g.zA; g.B += zb * g.zB; g.C += zb * g.zC;
  if (g.bias) g.bias += zb * g.zbias;
  if (g.frame_bias) g.frame_bias += zb * g.zfb;
  if (g.relu_mask) g.relu_mask += zb * g.zmask;
  const int kbeg = ks * g.k_per_split;
  const int kend = min(g.K, kbeg + g.k_per_split);
  const int tx = tid & 15, ty = tid >> 4;   // 16 x 16 threads, each 8 (m) x 8 (n) as 2x2 groups of 4x4

  // accumulators as packed FP32x2 pairs along n: one FFMA2 (a_i broadcast) updates two of them, i.e. half the
  // issue slots of scalar FFMA for the same FMA-pipe work - each half is an IEEE fma, results are unchanged
  f32x2 acc2[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc2[i][j] = pack2(0.f, 0.f);

  float ra[4], rb[4];
  TileLoader<AMODE> la;
  TileLoader<BMODE> lb;
  la.init(g.A, g.sam, g.sak, m0, g.M, kbeg, tid);
  lb.init(g.B, g.sbn, g.sbk, n0, g.N, kbeg, tid);
  la.load(kbeg, kend, ra);
  lb.load(kbeg, kend, rb);
  store_tile_smem<AMODE>(As[0], ra, tid);
  store_tile_smem<BMODE>(Bs[0], rb, tid);
  __syncthreads();

  int buf = 0;
  for (int k0 = kbeg; k0 < kend; k0 += GB_K) {
    const bool more = k0 + GB_K < kend;
    if (more) {
      la.load(k0 + GB_K, kend, ra);
      lb.load(k0 + GB_K, kend, rb);
    }
#pragma unroll
    for (int kk = 0; kk < GB_K; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const f32x2 b2[4] = {pack2(b0.x, b0.y), pack2(b0.z, b0.w), pack2(b1.x, b1.y), pack2(b1.z, b1.w)};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const f32x2 ai = pack2(a[i], a[i]);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc2[i][j] = fma2(ai, b2[j], acc2[i][j]);
      }
    }
    if (more) {
      store_tile_smem<AMODE>(As[buf ^ 1], ra, tid);
      store_tile_smem<BMODE>(Bs[buf ^ 1], rb, tid);
      __syncthreads();
      buf ^= 1;
    }
  }

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) unpack2(acc2[i][j], acc[i][2 * j], acc[i][2 * j + 1]);
  const bool split = g.splits > 1;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= g.M) continue;
    const float *fb = g.frame_bias ? g.frame_bias + (long long)(m / g.rows_per_frame) * g.ldfb : nullptr;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int nb = n0 + jh * 64 + tx * 4;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = nb + j;
        float x = acc[i][jh * 4 + j];
        if (n < g.N && (!split || ks == 0)) {
          if (g.bias) x += g.bias[n];
          if (fb) x += fb[n];
        }
        if (g.accumulate && !split && n < g.N) x += g.C[(long long)m * g.ldc + n];
        if (g.relu) x = fmaxf(x, 0.f);
        if (g.relu_mask && n < g.N) x = g.relu_mask[(long long)m * g.ldmask + n] > 0.f ? x : 0.f;
        v[j] = x;
      }
      float *crow = g.C + (long long)m * g.ldc + nb;
      if (split) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (nb + j < g.N) atomicAdd(crow + j, v[j]);
      } else if (nb + 3 < g.N && ((reinterpret_cast<uintptr_t>(crow) & 15) == 0)) {
        *reinterpret_cast<float4 *>(crow) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (nb + j < g.N) crow[j] = v[j];
      }
    }
  }
}

static int pick_mode(const float *p, long long s_mn, long long s_k) {
  const bool al = (reinterpret_cast<uintptr_t>(p) & 15) == 0;
  if (s_k == 1 && al && (s_mn % 4) == 0) return LOAD_KVEC;
  if (s_mn == 1 && al && (s_k % 4) == 0) return LOAD_MVEC;
  return LOAD_SCALAR;
}

template <int AM>
static void launch_b(int bm, const GemmArgs &g, dim3 grid, cudaStream_t st) {
  if (bm == LOAD_KVEC) gemm_f32_kernel<AM, LOAD_KVEC><<<grid, G_THREADS, 0, st>>>(g);
  else if (bm == LOAD_MVEC) gemm_f32_kernel<AM, LOAD_MVEC><<<grid, G_THREADS, 0, st>>>(g);
  else gemm_f32_kernel<AM, LOAD_SCALAR><<<grid, G_THREADS, 0, st>>>(g);
}

// split_k > 1 accumulates with atomics into C (C must hold the value to accumulate onto; relu and
// relu_mask are not allowed then).
int gemm_f32(GemmArgs g, int split_k, cudaStream_t st) {
  if (g.M <= 0 || g.N <= 0) return 0;
  if (g.K <= 0) { set_error("gemm_f32: K must be positive"); return PLR2_E_SHAPE; }
  if (split_k < 1) split_k = 1;
  if (split_k > 1 && (g.relu || g.relu_mask || !g.accumulate)) {
    set_error("gemm_f32: split-K needs accumulate=1 and a linear epilogue");
    return PLR2_E_UNSUPPORTED;
  }
  int kps = (g.K + split_k - 1) / split_k;
  kps = (kps + GB_K - 1) / GB_K * GB_K;
  split_k = (g.K + kps - 1) / kps;
  g.k_per_split = kps;
  g.splits = split_k;
  if (g.batch < 1) g.batch = 1;
  dim3 grid((g.N + GB_N - 1) / GB_N, (g.M + GB_M - 1) / GB_M, split_k * g.batch);
  if (grid.y > 65535) { set_error("gemm_f32: M too large (%d)", g.M); return PLR2_E_SHAPE; }
  const int am = pick_mode(g.A, g.sam, g.sak), bm = pick_mode(g.B, g.sbn, g.sbk);
  if (am == LOAD_KVEC) launch_b<LOAD_KVEC>(bm, g, grid, st);
  else if (am == LOAD_MVEC) launch_b<LOAD_MVEC>(bm, g, grid, st);
  else launch_b<LOAD_SCALAR>(bm, g, grid, st);
  return check_launch("gemm_f32");
}

}  // namespace plr2

extern "C" int plr2_gemm_f32(const plr2_gemm_desc *d, void *stream) {
  using namespace plr2;
  if (!d || !d->A || !d->B || !d->C) { set_error("plr2_gemm_f32: null pointer"); return PLR2_E_NULLPTR; }
  GemmArgs g;
  g.A = d->A; g.sam = d->sam; g.sak = d->sak;
  g.B = d->B; g.sbk = d->sbk; g.sbn = d->sbn;
  g.C = d->C; g.ldc = d->ldc;
  g.M = d->M; g.N = d->N; g.K = d->K;
  g.bias = d->bias;
  g.frame_bias = d->frame_bias; g.rows_per_frame = d->rows_per_frame > 0 ? d->rows_per_frame : 1; g.ldfb = d->ldfb;
  g.relu_mask = d->relu_mask; g.ldmask = d->ldmask;
  g.relu = d->relu; g.accumulate = d->accumulate;
  return gemm_f32(g, d->split_k, (cudaStream_t)stream);
}
