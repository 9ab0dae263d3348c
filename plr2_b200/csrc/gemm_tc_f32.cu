// Training-path contractions on the tcgen05 GEMM (gemm_bf16x2.cu).
//
// The forward/backward orchestration (posenet.cu) states every Conv1d(k=1)/Linear contraction of
// lib/network.py:54-63,107-121,152-163,193-200 and of its autograd as a GemmArgs problem on FP32
// operands in one of three layouts:
//     forward      Y  = X  W^T        A k-contiguous,  B k-contiguous      (gemm_nt)
//     input grad   dX = dY W          A k-contiguous,  B n-contiguous      (gemm_nn)
//     weight grad  dW = dY^T X        A m-contiguous,  B n-contiguous      (gemm_tn, k = points)
// The tensor-core kernel wants both operands K-major as split-bf16 planes (hi = bf16(v),
// lo = bf16(v - hi)), K a multiple of 64.  gemm_tc_f32 therefore packs A and B into scratch planes
// (a straight or a transposing split, zero-padded in K; both operands in one launch), runs ONE gemm_bf16x2 launch (3 MMAs per
// MAC, FP32 accumulation, FP32 output through TMA stores) and, where the problem accumulates onto C
// or passes through a ReLU mask, finishes with an element-wise combine out of a scratch C.
// The packs cost 8 bytes per operand element of HBM traffic; at the training shapes (8 frames =
// 8000 rows) every problem is launch/latency bound either way.
#include <stdlib.h>
#include "gemm.cuh"
#include "gemm_tc.cuh"

namespace plr2 {

constexpr int PACK_TILE = 32;

// One operand of a pack launch:
//   dst[plane p][z][r][kk] (row pitch ldd, chunk stride dst_z) = split(src[r * s_row + k * s_col]),
//   k = z * cols_z + kk for k < cols, 0 beyond; z walks the K chunks of a split-K problem (one chunk of
//   cols_z = padded K columns otherwise).
// trans (src contiguous along r): 32 x 32 tiles go through shared memory so that both the reads (along r)
// and the writes (along k) are coalesced.  Otherwise a tile is 32 rows x 128 columns, four consecutive k per
// thread: one float4 load where the source allows it (vec), two 8-byte stores.
struct PackDesc {
  const float *src; long long s_row, s_col;
  int rows, cols, cols_z, chunks;
  __nv_bfloat16 *dst; long long ldd, plane, dst_z;
  int trans, vec;
  int tiles_x, tiles_y, blocks;
  float *colsum; long long colsum_ld;   // straight packs only: colsum[row tile][column] = sum of the tile's 32 rows
};

__device__ __forceinline__ void split_store4(__nv_bfloat16 *hi_p, __nv_bfloat16 *lo_p, const float (&v)[4]) {
  __nv_bfloat16 h[4], l[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    h[e] = __float2bfloat16_rn(v[e]);
    l[e] = __float2bfloat16_rn(v[e] - __bfloat162float(h[e]));
  }
  *reinterpret_cast<uint2 *>(hi_p) = *reinterpret_cast<const uint2 *>(h);
  *reinterpret_cast<uint2 *>(lo_p) = *reinterpret_cast<const uint2 *>(l);
}

__device__ __forceinline__ void pack_tile(const PackDesc &d, int local, float (*tile)[PACK_TILE + 1]) {
  const int per_z = d.tiles_x * d.tiles_y;
  const int z = local / per_z, t = local - z * per_z;
  const int bx = t % d.tiles_x, by = t / d.tiles_x;
  const int kz = z * d.cols_z;                       // first source column of this chunk
  const int tx = threadIdx.x, ty = threadIdx.y;
  __nv_bfloat16 *dst = d.dst + z * d.dst_z;
  const int r0 = by * PACK_TILE;
  if (d.trans) {
    const int k0 = bx * PACK_TILE;
#pragma unroll
    for (int j = 0; j < PACK_TILE / 8; ++j) {
      const int kk = k0 + ty + 8 * j, r = r0 + tx;
      tile[ty + 8 * j][tx] = (kk < d.cols_z && kz + kk < d.cols && r < d.rows)
                                 ? __ldg(d.src + (long long)r * d.s_row + (long long)(kz + kk) * d.s_col) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < PACK_TILE / 8; ++j) {
      const int r = r0 + ty + 8 * j, kk = k0 + tx;
      if (r < d.rows && kk < d.cols_z) {
        const float v = tile[tx][ty + 8 * j];
        const __nv_bfloat16 hi = __float2bfloat16_rn(v);
        dst[(long long)r * d.ldd + kk] = hi;
        dst[d.plane + (long long)r * d.ldd + kk] = __float2bfloat16_rn(v - __bfloat162float(hi));
      }
    }
    return;
  }
  const int kk = bx * (4 * PACK_TILE) + 4 * tx;      // cols_z is a multiple of 64: kk < cols_z covers kk..kk+3
  float cs[4] = {0.f, 0.f, 0.f, 0.f};                // column sums over this thread's rows (fixed order)
#pragma unroll
  for (int j = 0; j < PACK_TILE / 8; ++j) {
    const int r = r0 + ty + 8 * j;
    if (r < d.rows && kk < d.cols_z) {
      float v[4];
      const float *sp = d.src + (long long)r * d.s_row + (long long)(kz + kk) * d.s_col;
      if (d.vec && kz + kk + 3 < d.cols) {
        const float4 q = __ldg(reinterpret_cast<const float4 *>(sp));
        v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = kz + kk + e < d.cols ? __ldg(sp + e * d.s_col) : 0.f;
      }
      __nv_bfloat16 *o = dst + (long long)r * d.ldd + kk;
      split_store4(o, o + d.plane, v);
#pragma unroll
      for (int e = 0; e < 4; ++e) cs[e] += v[e];
    }
  }
  if (d.colsum != nullptr) {                          // block-uniform; the tile array is free on this path
    float *part = &tile[0][0];                        // [8 row lanes][128 columns]
#pragma unroll
    for (int e = 0; e < 4; ++e) part[ty * (4 * PACK_TILE) + 4 * tx + e] = cs[e];
    __syncthreads();
    const int t = ty * PACK_TILE + tx;                // 256 threads, 128 columns: the lower half sums
    if (t < 4 * PACK_TILE) {
      const int c = bx * (4 * PACK_TILE) + t;
      if (c < d.cols) {
        float a = 0.f;
#pragma unroll
        for (int l = 0; l < 8; ++l) a += part[l * (4 * PACK_TILE) + t];
        d.colsum[(long long)by * d.colsum_ld + c] = a;
      }
    }
  }
}

// Both operands of a contraction in one launch: blocks [0, a.blocks) pack A, the rest pack B.
__global__ void __launch_bounds__(PACK_TILE * 8)
pack_split_kernel(PackDesc a, PackDesc b) {
  __shared__ float tile[PACK_TILE][PACK_TILE + 1];
  const int bid = blockIdx.x;
  if (bid < a.blocks) pack_tile(a, bid, tile);
  else pack_tile(b, bid - a.blocks, tile);
}

static PackDesc pack_desc(const float *src, long long s_row, long long s_col, int rows, int cols, int cols_z, int chunks,
                          __nv_bfloat16 *dst, long long ldd, long long plane, long long dst_z) {
  PackDesc d;
  d.src = src; d.s_row = s_row; d.s_col = s_col; d.rows = rows; d.cols = cols; d.cols_z = cols_z; d.chunks = chunks;
  d.dst = dst; d.ldd = ldd; d.plane = plane; d.dst_z = dst_z;
  d.trans = (s_row == 1 && s_col != 1) ? 1 : 0;
  // float4 loads: rows, chunks and columns must start on 16-byte boundaries of the source
  d.vec = (!d.trans && s_col == 1 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 && s_row % 4 == 0 && cols_z % 4 == 0) ? 1 : 0;
  d.tiles_x = d.trans ? (cols_z + PACK_TILE - 1) / PACK_TILE : (cols_z + 4 * PACK_TILE - 1) / (4 * PACK_TILE);
  d.tiles_y = (rows + PACK_TILE - 1) / PACK_TILE;
  d.blocks = d.tiles_x * d.tiles_y * chunks;
  d.colsum = nullptr; d.colsum_ld = 0;
  return d;
}

// What the A scratch planes currently hold (per host thread: a forward/backward is issued by one thread).
struct PlaneRecord {
  const float *src = nullptr; long long s_row = 0; int rows = 0, cols = 0; long long pitch = 0;
  const __nv_bfloat16 *dst = nullptr; cudaStream_t st = nullptr;
};
static thread_local PlaneRecord g_a_planes;

static int pack_pair(const PackDesc &a, const PackDesc &b, cudaStream_t st) {
  const long long blocks = (long long)a.blocks + b.blocks;
  if (blocks <= 0 || blocks > 0x7fffffffLL) { set_error("pack_split: bad block count"); return PLR2_E_SHAPE; }
  pack_split_kernel<<<(unsigned)blocks, dim3(PACK_TILE, 8), 0, st>>>(a, b);
  return check_launch("pack_split");
}

// C = ((accumulate ? C : 0) + sum_z Ct[:, z*N + n]) * (mask > 0)   (the epilogue order of gemm_f32_kernel;
// nsum > 1: partial products of a split-K problem, summed in chunk order)
// VEC: four consecutive columns per thread through 16-byte accesses (N, the pitches and the bases allow it).
template <bool VEC>
__global__ void __launch_bounds__(256)
combine_kernel(const float *__restrict__ ct, long long ldct, int nsum, float *C, long long ldc, const float *mask,
               long long ldmask, int M, int N, int accumulate) {
  constexpr int W = VEC ? 4 : 1;
  const unsigned ncw = (unsigned)(N / W);
  const unsigned total = (unsigned)M * ncw;                    // < 2^31 (checked by the launcher)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned r = i / ncw;
    const int n = (int)(i - r * ncw) * W;
    const float *src = ct + r * ldct + n;
    float *c = C + r * ldc + n;
    float v[W];
    if (VEC) {
      float4 q = __ldg(reinterpret_cast<const float4 *>(src));
      for (int z = 1; z < nsum; ++z) {
        const float4 p = __ldg(reinterpret_cast<const float4 *>(src + (long long)z * N));
        q.x += p.x; q.y += p.y; q.z += p.z; q.w += p.w;
      }
      v[0] = q.x; v[1 % W] = q.y; v[2 % W] = q.z; v[3 % W] = q.w;
      if (accumulate) {
        const float4 p = *reinterpret_cast<const float4 *>(c);
        v[0] += p.x; v[1 % W] += p.y; v[2 % W] += p.z; v[3 % W] += p.w;
      }
      if (mask != nullptr) {                                   // mask may alias C: same thread, same elements
        const float4 m = *reinterpret_cast<const float4 *>(mask + r * ldmask + n);
        if (!(m.x > 0.f)) v[0] = 0.f;
        if (!(m.y > 0.f)) v[1 % W] = 0.f;
        if (!(m.z > 0.f)) v[2 % W] = 0.f;
        if (!(m.w > 0.f)) v[3 % W] = 0.f;
      }
      *reinterpret_cast<float4 *>(c) = make_float4(v[0], v[1 % W], v[2 % W], v[3 % W]);
    } else {
      v[0] = src[0];
      for (int z = 1; z < nsum; ++z) v[0] += src[(long long)z * N];
      if (accumulate) v[0] += *c;
      if (mask != nullptr && !(mask[r * ldmask + n] > 0.f)) v[0] = 0.f;
      *c = v[0];
    }
  }
}

static int combine(const float *ct, long long ldct, int nsum, float *C, long long ldc, const float *mask, long long ldmask,
                   int M, int N, int accumulate, cudaStream_t st) {
  if ((long long)M * N >= 0x7fffffffLL) { set_error("gemm_tc_f32: combine: %d x %d elements", M, N); return PLR2_E_SHAPE; }
  auto al = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool vec = N % 4 == 0 && ldct % 4 == 0 && ldc % 4 == 0 && al(ct) && al(C) && (mask == nullptr || (ldmask % 4 == 0 && al(mask)));
  const long long items = (long long)M * (vec ? N / 4 : N);
  const int blocks = (int)((items + 255) / 256 < 148 * 8 ? (items + 255) / 256 : 148 * 8);
  if (vec) combine_kernel<true><<<blocks, 256, 0, st>>>(ct, ldct, nsum, C, ldc, mask, ldmask, M, N, accumulate);
  else combine_kernel<false><<<blocks, 256, 0, st>>>(ct, ldct, nsum, C, ldc, mask, ldmask, M, N, accumulate);
  return 0;
}

// ---- sign-safe ReLU of a training forward --------------------------------------------------
// The ReLU output of a training forward is the mask of its backward, and the split-bf16 product carries a
// relative error of up to 2^-16 per operand and 2^-16 for the dropped lo*lo term: a pre-activation closer to
// zero than SIGN_TAU * ||x_r|| * ||w_n|| (Cauchy-Schwarz over the K terms; SIGN_TAU = 2^-14 > 3 * 2^-16 plus the
// FP32 accumulation) may have the wrong sign.  Those elements - about one in a thousand - are recomputed as a
// sequential FP32 dot product before the ReLU; everywhere else the tensor-core sign is provably the true one.
constexpr float SIGN_TAU = 6.103515625e-05f;   // 2^-14

// out[z][r] = || src_z[r, 0:cols] ||_2, one warp per row (rows contiguous along k for both operands of a forward).
// One launch serves both operands: blocks [0, a.blocks) take the activations' rows, the rest the weights' rows.
struct NormDesc { const float *src; long long s_row, s_col, z_stride; int rows, cols, blocks; float *out; };
__global__ void __launch_bounds__(256)
row_norm_kernel(NormDesc a, NormDesc b) {
  const bool second = (int)blockIdx.x >= a.blocks;
  const NormDesc &d = second ? b : a;
  const int bid = second ? blockIdx.x - a.blocks : blockIdx.x;
  const int warp = (bid * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= d.rows) return;
  const float *p = d.src + blockIdx.y * d.z_stride + (long long)warp * d.s_row;
  float acc = 0.f;
  for (int k = lane; k < d.cols; k += 32) { const float v = __ldg(p + (long long)k * d.s_col); acc = fmaf(v, v, acc); }
  acc = warp_sum(acc);
  if (lane == 0) d.out[(size_t)blockIdx.y * d.rows + warp] = sqrtf(acc);
}

// C (pre-activation, written by the tensor-core launch without ReLU) -> relu(C), with the elements inside the
// error bound recomputed from the FP32 operands.  blockIdx.y = problem z of a batched launch.
// W = 4: four consecutive columns per lane through 16-byte accesses (N, ldc and the bases allow it).
template <int W>
__global__ void __launch_bounds__(256)
relu_signsafe_kernel(GemmArgs g, const float *__restrict__ nx, const float *__restrict__ nw) {
  const int z = blockIdx.y;
  const float *A = g.A + z * g.zA, *B = g.B + z * g.zB;
  float *C = g.C + z * g.zC;
  const float *bias = g.bias ? g.bias + z * g.zbias : nullptr;
  const float *fb = g.frame_bias ? g.frame_bias + z * g.zfb : nullptr;
  nx += (size_t)z * g.M;
  nw += (size_t)z * g.N;
  const unsigned ncw = (unsigned)(g.N / W);
  const unsigned total = (unsigned)g.M * ncw;                  // < 2^31 (checked by the launcher)
  const unsigned lane = threadIdx.x & 31;
  // warp-uniform trip count: every lane stays in the loop so that flagged elements can be recomputed by the
  // whole warp (coalesced reads along k, butterfly sum) instead of by one lane walking K serially
  for (unsigned base = (blockIdx.x * blockDim.x + threadIdx.x) - lane; base < total; base += gridDim.x * blockDim.x) {
    const unsigned i = base + lane;
    const bool live = i < total;
    const int r = live ? (int)(i / ncw) : 0;
    const int n = live ? (int)(i - (unsigned)r * ncw) * W : 0;
    float *cp = C + (long long)r * g.ldc + n;
    float y[W];
    if (W == 4) {
      const float4 q = live ? *reinterpret_cast<const float4 *>(cp) : make_float4(1.f, 1.f, 1.f, 1.f);
      y[0] = q.x; y[1 % W] = q.y; y[2 % W] = q.z; y[3 % W] = q.w;
    } else {
      y[0] = live ? *cp : 1.f;
    }
    const float bound = live ? SIGN_TAU * nx[r] : 0.f;
#pragma unroll
    for (int j = 0; j < W; ++j) {
      const bool flag = live && fabsf(y[j]) < bound * nw[n + j];
      unsigned todo = __ballot_sync(0xffffffffu, flag);
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const int rr = __shfl_sync(0xffffffffu, r, src), nn = __shfl_sync(0xffffffffu, n, src) + j;
        const float *a = A + (long long)rr * g.sam, *b = B + (long long)nn * g.sbn;
        float acc = 0.f;
        for (int k = lane; k < g.K; k += 32) acc = fmaf(__ldg(a + (long long)k * g.sak), __ldg(b + (long long)k * g.sbk), acc);
        acc = warp_sum(acc);
        if ((int)lane == src) {
          if (bias) acc += bias[nn];
          if (fb) acc += fb[(long long)(rr / g.rows_per_frame) * g.ldfb + nn];
          y[j] = acc;
        }
      }
    }
    if (live) {
      if (W == 4) *reinterpret_cast<float4 *>(cp) = make_float4(fmaxf(y[0], 0.f), fmaxf(y[1 % W], 0.f), fmaxf(y[2 % W], 0.f), fmaxf(y[3 % W], 0.f));
      else *cp = fmaxf(y[0], 0.f);
    }
  }
}

// PLR2_TC_WGRAD_MN=0 restores the transposing packs for the weight gradients (A/B experiments)
static bool tc_wgrad_mn() {
  static int v = -1;
  if (v < 0) { const char *e = getenv("PLR2_TC_WGRAD_MN"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}

bool gemm_tc_eligible(const GemmArgs &g) {
  // tiny N (conv1's 3 input channels as a weight-gradient N) and K stay on the SIMT kernel
  return g.M > 0 && g.N >= 16 && g.K >= 8 && !(g.relu && g.accumulate);
}

// Weight-gradient layout (A m-contiguous, B n-contiguous: both operands are [k rows][columns] in memory): the
// tensor cores read such planes MN-major, so the split is a straight element-wise pass (float4 in, 8-byte stores
// out, no transpose through shared memory, no K padding - rows past K read as zero through the tensor map).
// `heads` > 1: stacked problems whose dY / X are adjacent column blocks of ONE matrix each (the r|t|c heads): the
// matrices are split once, full width, and every head runs on its column range of the planes.
// Returns -1 when the scratch cannot hold the problem (the caller falls back to the transposing path).
static int wgrad_mn(const GemmArgs &g, int heads, const TcScratch &S, cudaStream_t st) {
  const int Mt = g.M * heads, Nt = g.N * heads;
  const int Mp = (Mt + 63) / 64 * 64, Np = (Nt + 63) / 64 * 64;
  if ((size_t)g.K * Mp > S.elems || (size_t)g.K * Np > S.elems) return -1;
  const long long tiles = (long long)((g.M + 255) / 256) * ((g.N + 127) / 128) * heads;
  long long want = (74 + tiles - 1) / tiles;
  if (want > g.K / 512) want = g.K / 512;
  if (want > 32) want = 32;
  while (want > 1 && (size_t)Mt * want * g.N > S.elems) --want;
  int splits = (int)(want < 1 ? 1 : want);
  const int Ks = ((g.K + splits - 1) / splits + 63) / 64 * 64;
  splits = (g.K + Ks - 1) / Ks;
  const bool need_ct = splits > 1 || g.accumulate || g.relu_mask != nullptr;
  if (need_ct && (size_t)Mt * splits * g.N > S.elems) return -1;
  int rc;
  PackDesc da = pack_desc(g.A, g.sak, 1, g.K, Mt, Mp, 1, S.pa, Mp, (long long)S.elems, 0);
  if (g.colsum_part != nullptr && (long long)da.tiles_y * Mt <= g.colsum_cap) {
    // the bias gradient's first stage rides on the split of dY: per 32-row tile column sums, summed by the caller
    da.colsum = g.colsum_part; da.colsum_ld = Mt;
    if (g.colsum_tiles) *g.colsum_tiles = da.tiles_y;
  }
  if ((rc = pack_pair(da, pack_desc(g.B, g.sbk, 1, g.K, Nt, Np, 1, S.pb, Np, (long long)S.elems, 0), st))) return rc;
  // dY as [rows][Mp] planes is also the K-major A operand of the input gradient that follows
  g_a_planes.src = g.A; g_a_planes.s_row = g.sak; g_a_planes.rows = g.K; g_a_planes.cols = Mt;
  g_a_planes.pitch = Mp; g_a_planes.dst = S.pa; g_a_planes.st = st;
  for (int z = 0; z < heads; ++z) {
    float *ct = S.ct + (size_t)z * g.M * splits * g.N;
    float *C = g.C + z * g.zC;
    TcGemm t;
    t.A = S.pa + (size_t)z * g.M; t.lda = Mp; t.a_plane = (long long)S.elems;
    t.W = S.pb + (size_t)z * g.N; t.ldw = Np; t.w_plane = (long long)S.elems;
    t.M = g.M; t.N = g.N; t.K = Ks; t.batch = splits; t.a_col_z = Ks; t.b_row_z = 0;
    t.mn_major = 1; t.k_rows = g.K;
    t.out_mode = TC_OUT_F32;
    if (need_ct) { t.Cf = ct; t.ldc = (long long)splits * g.N; t.c_col_z = g.N; }
    else { t.Cf = C; t.ldc = g.ldc; t.c_col_z = 0; }
    if ((rc = gemm_bf16x2(t, st))) return rc;
    if (need_ct) {
      if ((rc = combine(ct, (long long)splits * g.N, splits, C, g.ldc, g.relu_mask ? g.relu_mask + z * g.zmask : nullptr,
                        g.ldmask, g.M, g.N, g.accumulate, st))) return rc;
    }
  }
  return need_ct ? check_launch("gemm_tc_f32: combine", heads) : 0;
}

int gemm_tc_f32(const GemmArgs &g, const TcScratch &S, cudaStream_t st) {
  const int batch = g.batch < 1 ? 1 : g.batch;
  const int Kp = (g.K + 63) / 64 * 64;
  if (g.sam == 1 && g.sbn == 1 && !g.relu && g.bias == nullptr && g.frame_bias == nullptr && tc_wgrad_mn() &&
      (batch == 1 || (g.zA == g.M && g.zB == g.N && g.M % 8 == 0 && g.N % 8 == 0))) {
    const int rc = wgrad_mn(g, batch, S, st);
    if (rc >= 0) return rc;
  }
  // One launch serves the whole batch when problem z differs by a COLUMN offset of C (the r/t/c heads of
  // the forward and of the input gradient); stacked weight gradients (z advances whole matrices) are looped.
  const bool one = batch == 1 || (g.frame_bias == nullptr && g.zC >= g.N && g.zC * (batch - 1) + g.N <= g.ldc && g.N % 16 == 0);
  if (!one) {
    for (int z = 0; z < batch; ++z) {
      GemmArgs gz = g;
      gz.batch = 1;
      gz.A += z * g.zA; gz.B += z * g.zB; gz.C += z * g.zC;
      if (gz.bias) gz.bias += z * g.zbias;
      if (gz.frame_bias) gz.frame_bias += z * g.zfb;
      if (gz.relu_mask) gz.relu_mask += z * g.zmask;
      const int rc = gemm_tc_f32(gz, S, st);
      if (rc) return rc;
    }
    return 0;
  }
  int rc;
  // Split-K: a contraction over the points (weight gradients, K = frames * num_points) with a handful of
  // output tiles would run on a handful of SMs.  Its K chunks become the batch dimension of the launch
  // (chunk z = columns [z*Ks, +Ks) of the packed A, row block z of the packed B) and the partial products
  // meet in the combine pass.
  int splits = 1;
  if (batch == 1 && !g.relu && g.bias == nullptr && g.frame_bias == nullptr && g.K >= 1024 && g.N % 16 == 0) {
    const long long tiles = (long long)((g.M + 255) / 256) * ((g.N + 127) / 128);
    long long want = (74 + tiles - 1) / tiles;
    if (want > g.K / 512) want = g.K / 512;
    if (want > 32) want = 32;
    while (want > 1 && (size_t)g.M * want * g.N > S.elems) --want;
    splits = (int)(want < 1 ? 1 : want);
  }
  if (splits > 1) {
    const int Ks = ((g.K + splits - 1) / splits + 63) / 64 * 64;
    splits = (g.K + Ks - 1) / Ks;
    const size_t kp_all = (size_t)splits * Ks;
    if ((size_t)g.M * kp_all > S.elems || (size_t)g.N * kp_all > S.elems) return gemm_f32(g, 1, st);
    g_a_planes.src = nullptr;
    if ((rc = pack_pair(pack_desc(g.A, g.sam, g.sak, g.M, g.K, (int)kp_all, 1, S.pa, (long long)kp_all, (long long)S.elems, 0),
                        pack_desc(g.B, g.sbn, g.sbk, g.N, g.K, Ks, splits, S.pb, Ks, (long long)S.elems, (long long)g.N * Ks), st))) return rc;
    TcGemm t;
    t.A = S.pa; t.lda = (long long)kp_all; t.a_plane = (long long)S.elems;
    t.W = S.pb; t.ldw = Ks; t.w_plane = (long long)S.elems;
    t.M = g.M; t.N = g.N; t.K = Ks; t.batch = splits;
    t.a_col_z = Ks; t.b_row_z = g.N;
    t.out_mode = TC_OUT_F32; t.Cf = S.ct; t.ldc = (long long)splits * g.N; t.c_col_z = g.N;
    if ((rc = gemm_bf16x2(t, st))) return rc;
    if ((rc = combine(S.ct, (long long)splits * g.N, splits, g.C, g.ldc, g.relu_mask, g.ldmask, g.M, g.N, g.accumulate, st))) return rc;
    return check_launch("gemm_tc_f32: combine");
  }
  const bool need_ct = g.accumulate || g.relu_mask != nullptr;
  const size_t a_elems = (size_t)g.M * batch * Kp, b_elems = (size_t)g.N * batch * Kp, c_elems = (size_t)g.M * batch * g.N;
  if (a_elems > S.elems || b_elems > S.elems || (need_ct && c_elems > S.elems)) return gemm_f32(g, 1, st);
  const PlaneRecord &R = g_a_planes;
  // (batched: problem z reads columns [z*zA, +K) of ONE matrix; its planes line up when the blocks are unpadded)
  const bool reuse_a = g.reuse_a_planes && g.sak == 1 && R.src == g.A && R.s_row == g.sam && R.rows == g.M &&
                       R.cols == batch * g.K && R.pitch == (long long)batch * Kp && R.dst == S.pa && R.st == st &&
                       (batch == 1 || (g.zA == g.K && g.K == Kp));
  if (!reuse_a) g_a_planes.src = nullptr;
  for (int z = 0; z < batch; ++z) {
    PackDesc pa_d = pack_desc(g.A + z * g.zA, g.sam, g.sak, g.M, g.K, Kp, 1, S.pa + (size_t)z * Kp, (long long)batch * Kp, (long long)S.elems, 0);
    if (reuse_a) pa_d.blocks = 0;                     // the planes of the weight gradient before are still in place
    if ((rc = pack_pair(pa_d,
                        pack_desc(g.B + z * g.zB, g.sbn, g.sbk, g.N, g.K, Kp, 1, S.pb + (size_t)z * g.N * Kp, Kp, (long long)S.elems, 0), st))) return rc;
  }
  TcGemm t;
  t.A = S.pa; t.lda = (long long)batch * Kp; t.a_plane = (long long)S.elems;
  t.W = S.pb; t.ldw = Kp; t.w_plane = (long long)S.elems;
  t.M = g.M; t.N = g.N; t.K = Kp; t.batch = batch;
  t.a_col_z = Kp; t.b_row_z = g.N;
  t.bias = g.bias; t.bias_z = g.zbias;
  t.frame_bias = g.frame_bias; t.rows_per_frame = g.rows_per_frame; t.ldfb = g.ldfb;
  // ReLU layers (training forwards) go through the sign-safe pass; it borrows the unused C staging area for the
  // row norms of both operands
  const bool sign_safe = g.relu && !need_ct && (size_t)batch * ((size_t)g.M + g.N) <= S.elems;
  t.relu = sign_safe ? 0 : g.relu; t.out_mode = TC_OUT_F32;
  if (need_ct) { t.Cf = S.ct; t.ldc = (long long)batch * g.N; t.c_col_z = g.N; }
  else { t.Cf = g.C; t.ldc = g.ldc; t.c_col_z = g.zC; }
  if ((rc = gemm_bf16x2(t, st))) return rc;
  if (sign_safe) {
    float *nx = S.ct, *nw = S.ct + (size_t)batch * g.M;
    const NormDesc na{g.A, g.sam, g.sak, g.zA, g.M, g.K, (g.M + 7) / 8, nx}, nb{g.B, g.sbn, g.sbk, g.zB, g.N, g.K, (g.N + 7) / 8, nw};
    row_norm_kernel<<<dim3(na.blocks + nb.blocks, batch), 256, 0, st>>>(na, nb);
    if ((long long)g.M * g.N >= 0x7fffffffLL) { set_error("gemm_tc_f32: sign-safe relu: %d x %d elements", g.M, g.N); return PLR2_E_SHAPE; }
    const bool vec = g.N % 4 == 0 && g.ldc % 4 == 0 && g.zC % 4 == 0 && (reinterpret_cast<uintptr_t>(g.C) & 15) == 0;
    const long long items = (long long)g.M * (vec ? g.N / 4 : g.N);
    const int blocks = (int)((items + 255) / 256 < 148 * 16 ? (items + 255) / 256 : 148 * 16);
    GemmArgs gs = g;
    gs.batch = batch;
    if (vec) relu_signsafe_kernel<4><<<dim3(blocks, batch), 256, 0, st>>>(gs, nx, nw);
    else relu_signsafe_kernel<1><<<dim3(blocks, batch), 256, 0, st>>>(gs, nx, nw);
    if ((rc = check_launch("gemm_tc_f32: sign-safe relu", 2))) return rc;
  }
  if (need_ct) {
    for (int z = 0; z < batch; ++z) {
      if ((rc = combine(S.ct + (size_t)z * g.N, (long long)batch * g.N, 1, g.C + z * g.zC, g.ldc,
                        g.relu_mask ? g.relu_mask + z * g.zmask : nullptr, g.ldmask, g.M, g.N, g.accumulate, st))) return rc;
    }
    if ((rc = check_launch("gemm_tc_f32: combine", batch))) return rc;
  }
  return 0;
}

}  // namespace plr2
