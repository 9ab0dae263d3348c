// Kernel (2): the dense contractions of the PointNet / head MLPs on the 5th-gen tensor cores.
//
//   Y[M,N] = epi( X[M,K] W[N,K]^T )   with X, W given as SPLIT-BF16:  v = hi + lo, hi = bf16(v),
//   lo = bf16(v - hi)  (16 mantissa bits), and the product evaluated as
//        X_hi W_hi + X_hi W_lo + X_lo W_hi          (3 tcgen05.mma kind::f16, FP32 accumulate in TMEM)
// which keeps ~2^-16 relative operand error — enough for the north-star rtol 1e-3 on poses, losses and
// gradients, which plain TF32 / bf16 operands do not reach on this network.
//
// Structure (one CTA per SM, persistent over output tiles of 128 x BN):
//   warp 0    : TMA producer — per K-block of 64: X_hi, X_lo (128x64 bf16 each), W_hi, W_lo (BNx64)
//               into a STAGES-deep shared-memory ring (128B-swizzled, K-major), mbarrier expect_tx
//   warp 1    : MMA issuer — 3 products x 4 (UMMA_K = 16) tcgen05.mma per K-block into one of two
//               TMEM accumulator stages; tcgen05.commit frees the smem slot / publishes the tile
//   warp 2    : TMEM allocator
//   warps 4-7 : epilogue — tcgen05.ld 32 lanes x 32 columns, + bias (+ per-frame bias), ReLU, then
//               FP32 store | split-bf16 store (operand of the next layer) | per-frame pooling partials
//               (AvgPool1d fused: conv6's output is never written)
#include "gemm_tc.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdlib.h>
#include <vector>

namespace plr2 {

using namespace tc;

constexpr int TC_BM = 128, TC_BK = 64;
constexpr int TC_EPI_WARPS = 8;                        // two warpgroups: each TMEM lane quarter is served by 2 warps
constexpr int TC_THREADS = 128 + 32 * TC_EPI_WARPS;
constexpr int TC_EPI_COLS = 16;                        // columns per epilogue chunk
constexpr int TC_EPI_PITCH = 20;                       // floats per staged row (16 + 4 pad: conflict-free float4 stores)
constexpr int TC_EPI_FLOATS = 32 * TC_EPI_PITCH;        // per epilogue warp (fallback store path)
constexpr int TC_SLAB_BYTES = 2048;                    // one 32-row x 16-column chunk: fp32, or bf16 hi + lo
constexpr int TC_SLABS = 2;                            // slabs per epilogue warp (TMA store of one in flight while the next fills)
static_assert(TC_EPI_FLOATS * 4 <= TC_SLABS * TC_SLAB_BYTES, "fallback transpose slab fits");

// BN = output-tile width, NPROD = 3 (hi+lo activations) | 2 (hi-only activations: X_hi W_hi + X_hi W_lo) |
// 1 (single product X_hi W_hi: neither lo plane is loaded),
// CG = 1 (one CTA per tile of 128 rows) | 2 (CTA pair, cta_group::2: one 256-row tile per pair, each CTA
// stages its own 128 rows of X and HALF of the W tile, so the L2->smem traffic per FLOP drops by a third
// and a third pipeline stage fits).
template <int BN, int NPROD, int CG> struct TcCfg {
  static constexpr int A_BYTES = TC_BM * TC_BK * 2;            // one plane
  static constexpr int B_ROWS = BN / CG;                       // W rows staged by this CTA
  static constexpr int B_BYTES = B_ROWS * TC_BK * 2;
  static constexpr int A_PLANES = NPROD == 3 ? 2 : 1;
  static constexpr int B_PLANES = NPROD == 1 ? 1 : 2;
  static constexpr int STAGE_BYTES = A_PLANES * A_BYTES + B_PLANES * B_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;                      // two accumulator stages
  static constexpr int EPI_BYTES = TC_EPI_WARPS * TC_SLABS * TC_SLAB_BYTES;   // per-warp rings of output slabs
  static constexpr int BIAS_BYTES = 2 * BN * 4;                             // [frame a|b][BN] staged bias of the current tile
  static constexpr int HEAD_BYTES = BN == 128 ? (2 * 4 * 128 + 8) * 4 + 128 * 4 * 4 : 0;   // TC_OUT_HEAD: class weights + half-sum exchange
  static constexpr int FIXED_BYTES = 256 /*barriers*/ + EPI_BYTES + BIAS_BYTES + HEAD_BYTES;
  static constexpr int STAGES_FIT = (232448 - FIXED_BYTES) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_FIT > 8 ? 8 : STAGES_FIT;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + FIXED_BYTES;
  static_assert(STAGES >= 2, "pipeline needs two stages");
  static_assert(2 * STAGES + 4 <= 30, "barrier area");
};

struct TcKernelArgs {
  int M, N, K, batch;
  int m_tiles, n_tiles;          // m_tiles counts tiles of 128*CG rows
  int a_col_z, b_row_z;          // per-batch offsets: A K-coordinate (elements), W row
  long long c_col_z, bias_z;
  const float *bias, *frame_bias;
  int rows_per_frame; long long ldfb;
  int relu, out_mode;
  float *Cf; long long ldc;
  __nv_bfloat16 *Cb; long long ldcb, c_plane;
  float *pool;                   // [ceil(M/32)][2][N]
  const float *w4, *b4; const int64_t *obj; int num_obj;   // TC_OUT_HEAD
  float *pred_r, *pred_t, *pred_c;
  int mn_major;                  // operands stored [k rows][mn columns] (TcGemm::mn_major)
  int fp16;                      // operand planes (and TC_OUT_BF16X2 outputs) are IEEE fp16 hi/lo instead of bf16 hi/lo
  int tma_store;                 // tile outputs leave through TMA stores (maps tmC_*)
  int debug;                     // PLR2_TC_DEBUG (timing experiments only): 1 = no global stores, 2 = no epilogue work
};

// v = hi + lo with hi = round(v), lo = round(v - hi), two values packed per 32-bit word
__device__ __forceinline__ void split2_bf16(float a, float b, uint32_t &hi, uint32_t &lo) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  hi = *reinterpret_cast<const uint32_t *>(&h);
  const float h0 = __uint_as_float(hi << 16), h1 = __uint_as_float(hi & 0xffff0000u);
  const __nv_bfloat162 l = __floats2bfloat162_rn(a - h0, b - h1);
  lo = *reinterpret_cast<const uint32_t *>(&l);
}
__device__ __forceinline__ void split2_f16(float a, float b, uint32_t &hi, uint32_t &lo) {
  const __half2 h = __floats2half2_rn(a, b);
  hi = *reinterpret_cast<const uint32_t *>(&h);
  const float2 hf = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
  lo = *reinterpret_cast<const uint32_t *>(&l);
}

template <int BN, int NPROD, int CG>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_bf16x2_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                   const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo,
                   const __grid_constant__ CUtensorMap tmC_hi, const __grid_constant__ CUtensorMap tmC_lo,
                   const TcKernelArgs g) {
  using Cfg = TcCfg<BN, NPROD, CG>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // The 128B-swizzled operand tiles need 1024-byte alignment; the dynamic shared window of a kernel
  // without static shared memory starts on such a boundary (checked, not assumed).
  uint8_t *smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
  uint64_t *full_bar = bars;                       // [STAGES]  TMA -> MMA   (CG == 2: the leader's is used)
  uint64_t *empty_bar = bars + Cfg::STAGES;        // [STAGES]  MMA -> TMA
  uint64_t *tfull_bar = bars + 2 * Cfg::STAGES;    // [2]       MMA -> epilogue
  uint64_t *tempty_bar = tfull_bar + 2;            // [2]       epilogue -> MMA (CG == 2: the leader's is used)
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);

  // shuffle broadcast: tells the compiler the warp index is uniform, so TMA/MMA operands derived from it
  // go straight to uniform registers instead of through R2UR waterfall loops
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;      // 0 = leader (issues the MMAs)
  const int unit = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;       // tile-processing unit (CTA or pair)
  const int n_units = CG == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int num_kb = g.K / TC_BK;
  const int tiles_per_z = g.m_tiles * g.n_tiles;
  const int num_tiles = tiles_per_z * g.batch;
  // Tile of round `it` for this unit.  The slot rotates by one per round so that every unit sees every
  // n-tile index equally often (with a fixed slot, unit u would only ever meet n = (u + it*n_units) % n_tiles,
  // and a half-width last n-tile would unbalance the units); each round still covers a contiguous tile range.
  auto tile_of = [&](int it) { return it * n_units + (unit + it) % n_units; };
  // Width of the MMA for n-tile `nt`: a last tile with at most BN/2 valid columns runs at N = BN/2
  // (MN-major operands arrive in 64-column blocks: always full width)
  auto tile_nw = [&](int nt) { return (!g.mn_major && g.N - nt * BN <= BN / 2) ? BN / 2 : BN; };

  if (threadIdx.x == 0) {
    prefetch_tensormap(&tmA_hi); prefetch_tensormap(&tmA_lo);
    prefetch_tensormap(&tmB_hi); prefetch_tensormap(&tmB_lo);
    for (int s = 0; s < Cfg::STAGES; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar + s, 1); mbar_init(tempty_bar + s, CG * TC_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 2) {
    if (CG == 2) tmem_alloc_cg2(tmem_slot, Cfg::TMEM_COLS);
    else tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  }
  fence_before_sync();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs of a pair) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int it = 0, t; (t = tile_of(it)) < num_tiles; ++it) {
        const int z = t / tiles_per_z, r = t % tiles_per_z;
        const int nt = r % g.n_tiles;
        // CG == 2: each CTA supplies half of the MMA's N columns, i.e. W rows [n_tile + rank*nw/2, +nw/2)
        // (the box still loads B_ROWS rows; a half-width MMA reads the first nw/2 of them)
        const int m0 = (r / g.n_tiles) * (TC_BM * CG) + rank * TC_BM, n0 = nt * BN + (CG == 2 ? rank * (tile_nw(nt) / 2) : 0);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar + stage, phase ^ 1);
          uint8_t *sA = smem + stage * Cfg::STAGE_BYTES;
          uint8_t *sB = sA + Cfg::A_PLANES * Cfg::A_BYTES;
          const int ka = z * g.a_col_z + kb * TC_BK, kw = kb * TC_BK, rw = z * g.b_row_z + n0;
          if (g.mn_major) {
            // boxes of 64 mn-columns x 64 k-rows, one 8 KB block each: A columns [m0, +128), W columns [rw, +B_ROWS)
            const uint32_t bar = CG == 2 ? mapa_u32(smem_u32(full_bar + stage), 0) : 0u;
            if (CG == 2) { if (rank == 0) mbar_arrive_expect_tx(full_bar + stage, 2 * Cfg::STAGE_BYTES); }
            else mbar_arrive_expect_tx(full_bar + stage, Cfg::STAGE_BYTES);
            auto ld = [&](uint8_t *dst, const CUtensorMap *tm, int c, int r) {
              if (CG == 2) tma_load_2d_cg2(dst, tm, bar, c, r); else tma_load_2d(dst, tm, full_bar + stage, c, r);
            };
#pragma unroll
            for (int j = 0; j < TC_BM / 64; ++j) {
              ld(sA + j * 8192, &tmA_hi, m0 + 64 * j, ka);
              if (NPROD == 3) ld(sA + Cfg::A_BYTES + j * 8192, &tmA_lo, m0 + 64 * j, ka);
            }
#pragma unroll
            for (int j = 0; j < Cfg::B_ROWS / 64; ++j) {
              ld(sB + j * 8192, &tmB_hi, rw + 64 * j, ka);
              if (NPROD != 1) ld(sB + Cfg::B_BYTES + j * 8192, &tmB_lo, rw + 64 * j, ka);
            }
          } else if (CG == 2) {
            const uint32_t bar = mapa_u32(smem_u32(full_bar + stage), 0);      // the leader's barrier
            if (rank == 0) mbar_arrive_expect_tx(full_bar + stage, 2 * Cfg::STAGE_BYTES);
            tma_load_2d_cg2(sA, &tmA_hi, bar, ka, m0);
            if (NPROD == 3) tma_load_2d_cg2(sA + Cfg::A_BYTES, &tmA_lo, bar, ka, m0);
            tma_load_2d_cg2(sB, &tmB_hi, bar, kw, rw);
            if (NPROD != 1) tma_load_2d_cg2(sB + Cfg::B_BYTES, &tmB_lo, bar, kw, rw);
          } else {
            mbar_arrive_expect_tx(full_bar + stage, Cfg::STAGE_BYTES);
            tma_load_2d(sA, &tmA_hi, full_bar + stage, ka, m0);
            if (NPROD == 3) tma_load_2d(sA + Cfg::A_BYTES, &tmA_lo, full_bar + stage, ka, m0);
            tma_load_2d(sB, &tmB_hi, full_bar + stage, kw, rw);
            if (NPROD != 1) tma_load_2d(sB + Cfg::B_BYTES, &tmB_lo, full_bar + stage, kw, rw);
          }
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (lane == 0 && rank == 0) {
      // operand planes are bf16 (training scratch, stand-alone entry) or fp16 (inference activations / packed weights)
      const uint32_t major = g.mn_major ? (IDESC_A_MN | IDESC_B_MN) : 0u;
      const uint32_t idesc_full = (g.fp16 ? make_idesc_f16(TC_BM * CG, BN) : make_idesc_bf16(TC_BM * CG, BN)) | major;
      const uint32_t idesc_half = (g.fp16 ? make_idesc_f16(TC_BM * CG, BN / 2) : make_idesc_bf16(TC_BM * CG, BN / 2)) | major;
      const uint64_t k_adv = g.mn_major ? 2048 >> 4 : 32 >> 4;   // one MMA's 16 k: 16 rows of 128 B, or 32 B inside the atom
      int stage = 0; uint32_t phase = 0;
      int as = 0; uint32_t aphase = 0;
      for (int it = 0, t; (t = tile_of(it)) < num_tiles; ++it) {
        const uint32_t idesc = tile_nw((t % tiles_per_z) % g.n_tiles) == BN ? idesc_full : idesc_half;
        mbar_wait(tempty_bar + as, aphase ^ 1);          // epilogues (of both CTAs) have drained this accumulator
        fence_after_sync();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar + stage, phase);
          fence_after_sync();
          const uint32_t sA = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t sB = sA + Cfg::A_PLANES * Cfg::A_BYTES;
          const uint64_t a_hi = g.mn_major ? make_desc_mn128(sA) : make_desc_k128(sA);
          const uint64_t a_lo = g.mn_major ? make_desc_mn128(sA + Cfg::A_BYTES) : make_desc_k128(sA + Cfg::A_BYTES);
          const uint64_t b_hi = g.mn_major ? make_desc_mn128(sB) : make_desc_k128(sB);
          const uint64_t b_lo = g.mn_major ? make_desc_mn128(sB + Cfg::B_BYTES) : make_desc_k128(sB + Cfg::B_BYTES);
#pragma unroll
          for (int k = 0; k < TC_BK / 16; ++k) {          // small cross terms first, then hi*hi
            const uint64_t adv = static_cast<uint64_t>(k) * k_adv;
            if (CG == 2) {
              if (NPROD == 3) mma_f16_cg2(tmem_d, a_lo + adv, b_hi + adv, idesc, (kb | k) != 0);
              if (NPROD != 1) mma_f16_cg2(tmem_d, a_hi + adv, b_lo + adv, idesc, NPROD == 3 ? 1 : (kb | k) != 0);
              mma_f16_cg2(tmem_d, a_hi + adv, b_hi + adv, idesc, NPROD == 1 ? (kb | k) != 0 : 1);
            } else {
              if (NPROD == 3) mma_f16(tmem_d, a_lo + adv, b_hi + adv, idesc, (kb | k) != 0);
              if (NPROD != 1) mma_f16(tmem_d, a_hi + adv, b_lo + adv, idesc, NPROD == 3 ? 1 : (kb | k) != 0);
              mma_f16(tmem_d, a_hi + adv, b_hi + adv, idesc, NPROD == 1 ? (kb | k) != 0 : 1);
            }
          }
          if (CG == 2) mma_commit_cg2(empty_bar + stage, 3); else mma_commit(empty_bar + stage);   // smem slot reusable
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        if (CG == 2) mma_commit_cg2(tfull_bar + as, 3); else mma_commit(tfull_bar + as);           // accumulator complete
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int we = warp - 4;                              // 0..7
    const int wq = warp & 3;                              // TMEM lane quarter == warp % 4
    const int half = we >> 2;                             // which half of the column chunks this warpgroup takes
    const int te = threadIdx.x - 128;                     // 0..255 among epilogue threads
    uint8_t *slabs = smem + Cfg::STAGES * Cfg::STAGE_BYTES + 256 + we * (TC_SLABS * TC_SLAB_BYTES);
    float *sT = reinterpret_cast<float *>(slabs);         // fallback store path: [32][20] transpose slab
    float *sB = reinterpret_cast<float *>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + 256 + Cfg::EPI_BYTES);
    int as = 0; uint32_t aphase = 0;
    unsigned slab_i = 0;                                  // ring position, runs across tiles
    // bias (+ per-frame bias) of a tile for this thread's staging slots i = te, te + 256 (< 2*BN)
    // (the two addends stay in separate registers until they are staged, so nothing waits on the loads here)
    auto load_bias = [&](int t, float (&bv)[2], float (&fv)[2]) {
      const int z = t / tiles_per_z, r = t % tiles_per_z;
      const int m0 = (r / g.n_tiles) * (TC_BM * CG) + rank * TC_BM, n0 = (r % g.n_tiles) * BN;
      const float *bias = g.bias ? g.bias + z * g.bias_z : nullptr;
      const int fa = m0 / g.rows_per_frame;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i = te + u * 32 * TC_EPI_WARPS;
        const int sel = i / BN, c = i - sel * BN, n = n0 + c;
        bv[u] = 0.f; fv[u] = 0.f;
        if (i < 2 * BN && n < g.N) {
          if (bias) bv[u] = __ldg(bias + n);
          if (g.frame_bias) {
            const long long fr = fa + sel;
            if (fr * g.rows_per_frame < g.M) fv[u] = __ldg(g.frame_bias + fr * g.ldfb + n);
          }
        }
      }
    };
    // TC_OUT_HEAD: last-layer weights of head z for the classes of the tile's <= 2 frames, slots
    // i = te + u*256 of sW = [frame a|b][4 outputs][128] (+ 8 biases at the end)
    float *sW = sB + 2 * BN;
    float *sX = sW + 2 * 4 * 128 + 8;                     // [128 rows][4] partial dots of the upper column half
    auto load_headw = [&](int t, float (&wv)[5]) {
      const int z = t / tiles_per_z, r = t % tiles_per_z;
      const int m0 = (r / g.n_tiles) * (TC_BM * CG) + rank * TC_BM;
      const int fa = m0 / g.rows_per_frame;
      const int nout = z == 0 ? 4 : (z == 1 ? 3 : 1), base = z == 0 ? 0 : (z == 1 ? 4 * g.num_obj : 7 * g.num_obj);
#pragma unroll
      for (int u = 0; u < 5; ++u) {
        const int i = te + u * 32 * TC_EPI_WARPS;           // u == 4: the 8 biases
        const int sel = u < 4 ? i >> 9 : (te >> 2) & 1, j = u < 4 ? (i >> 7) & 3 : te & 3;
        long long fr = fa + sel;
        if (fr * g.rows_per_frame >= g.M) fr = fa;
        long long o = __ldg(g.obj + fr);
        o = o < 0 ? 0 : (o >= g.num_obj ? g.num_obj - 1 : o);
        float v = 0.f;
        if (j < nout) {
          if (u < 4) v = __ldg(g.w4 + (size_t)(base + o * nout + j) * 128 + (i & 127));
          else if (te < 8) v = __ldg(g.b4 + base + o * nout + j);
        }
        wv[u] = v;
      }
    };
    float bnext[2] = {0.f, 0.f}, fnext[2] = {0.f, 0.f}, wnext[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (tile_of(0) < num_tiles) {
      load_bias(tile_of(0), bnext, fnext);
      if (BN == 128 && g.out_mode == TC_OUT_HEAD) load_headw(tile_of(0), wnext);
    }
    for (int it = 0, t; (t = tile_of(it)) < num_tiles; ++it) {
      const int z = t / tiles_per_z, r = t % tiles_per_z;
      const int m0 = (r / g.n_tiles) * (TC_BM * CG) + rank * TC_BM, n0 = (r % g.n_tiles) * BN;
      const int nw_half = tile_nw(r % g.n_tiles) / 2;    // columns per warpgroup half in this tile
      // Stage bias[n] (+ frame_bias[frame][n]) of this tile's <= 2 frames: sB[sel][c], sel = row >= split_row.
      // The values were fetched while the previous tile was processed; the next tile's are requested now.
      const int fa = m0 / g.rows_per_frame;
      const int split_row = (fa + 1) * g.rows_per_frame;  // first row of the tile's second frame
      named_bar_sync(2, 32 * TC_EPI_WARPS);               // every warp has finished reading the previous tile's sB
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i = te + u * 32 * TC_EPI_WARPS;
        if (i < 2 * BN) sB[i] = bnext[u] + fnext[u];
      }
      if (BN == 128 && g.out_mode == TC_OUT_HEAD) {
#pragma unroll
        for (int u = 0; u < 4; ++u) sW[te + u * 32 * TC_EPI_WARPS] = wnext[u];
        if (te < 8) sW[2 * 4 * 128 + te] = wnext[4];
      }
      if (tile_of(it + 1) < num_tiles) {
        load_bias(tile_of(it + 1), bnext, fnext);
        if (BN == 128 && g.out_mode == TC_OUT_HEAD) load_headw(tile_of(it + 1), wnext);
      }
      named_bar_sync(1, 32 * TC_EPI_WARPS);
      mbar_wait(tfull_bar + as, aphase);
      fence_after_sync();
      const int rbase = m0 + wq * 32;                     // first row of this warp's 32-row slab
      if (g.out_mode == TC_OUT_POOL || g.out_mode == TC_OUT_HEAD || g.tma_store) {
        // ---- fast path: warpgroup `half` owns columns [half*BN/2, +BN/2) in chunks of 16; the TMEM load
        // of chunk i+1 is in flight while chunk i is processed; tile outputs leave through per-warp
        // shared-memory slabs and TMA stores (no LSU traffic), pooled sums through a halving butterfly.
        constexpr int NCH = BN / 2 / TC_EPI_COLS;
        const int cbase = half * nw_half;
        const uint32_t tbase = tmem_base + (static_cast<uint32_t>(wq * 32) << 16) + as * BN + cbase;
        const int row = rbase + lane;
        const bool second = row >= split_row;
        const float *bsel = sB + (second ? BN : 0) + cbase;
        uint32_t raw[2][16];
        float hacc[4] = {0.f, 0.f, 0.f, 0.f};             // TC_OUT_HEAD: this row's partial last-layer dots
        if (g.debug != 2) tmem_ld_32x16(tbase, raw[0]);
#pragma unroll
        for (int ci = 0; ci < NCH; ++ci) {
          const int c0 = ci * TC_EPI_COLS;
          if (c0 >= nw_half || n0 + cbase + c0 >= g.N || g.debug == 2) break;          // warp-uniform
          tmem_ld_wait();
          if (ci + 1 < NCH && c0 + TC_EPI_COLS < nw_half && n0 + cbase + c0 + TC_EPI_COLS < g.N) tmem_ld_32x16(tbase + c0 + TC_EPI_COLS, raw[(ci + 1) & 1]);
          float a[16];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 bb = *reinterpret_cast<const float4 *>(bsel + c0 + 4 * q);
            a[4 * q + 0] = __uint_as_float(raw[ci & 1][4 * q + 0]) + bb.x;
            a[4 * q + 1] = __uint_as_float(raw[ci & 1][4 * q + 1]) + bb.y;
            a[4 * q + 2] = __uint_as_float(raw[ci & 1][4 * q + 2]) + bb.z;
            a[4 * q + 3] = __uint_as_float(raw[ci & 1][4 * q + 3]) + bb.w;
          }
          if (g.relu) {
#pragma unroll
            for (int j = 0; j < 16; ++j) a[j] = fmaxf(a[j], 0.f);
          }
          const int ncol = z * (int)g.c_col_z + n0 + cbase + c0;      // first output column of this chunk
          if (g.out_mode == TC_OUT_POOL) {
            // per-column sums over this warp's 32 rows, split at the frame boundary if it falls inside
            const bool straddle = rbase < split_row && rbase + 31 >= split_row;       // warp-uniform
            const bool live = row < g.M;
            float va[16], vb[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float v = live ? a[j] : 0.f;
              va[j] = (straddle && second) ? 0.f : v;       // first frame present in the slab (or the only one)
              vb[j] = (straddle && second) ? v : 0.f;
            }
            auto butterfly = [&](float (&v)[16]) {
#pragma unroll
              for (int w = 8; w >= 1; w >>= 1) {            // lane bit (2w) selects which half it keeps
                const bool up = lane & (2 * w);
#pragma unroll
                for (int j = 0; j < w; ++j) {
                  const float send = up ? v[j] : v[j + w];
                  const float keep = up ? v[j + w] : v[j];
                  v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 2 * w);
                }
              }
              v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
            };
            butterfly(va);
            if (straddle) butterfly(vb);
            const int n = n0 + cbase + c0 + (lane >> 1);    // column held by this lane pair
            if ((lane & 1) == 0 && n < g.N && rbase < g.M && g.debug != 1) {
              // pool[(row group)][slot]: slot 0 = frame of the group's first row, slot 1 = the next frame
              float *dst = g.pool + ((long long)(rbase >> 5) * 2) * g.N + n;
              dst[0] = va[0];
              if (straddle) dst[g.N] = vb[0];             // slot 1 is read only for straddling groups
            }
          } else if (BN == 128 && g.out_mode == TC_OUT_HEAD) {
            const float *wsel = sW + (second ? 4 * 128 : 0) + cbase + c0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float4 w = *reinterpret_cast<const float4 *>(wsel + j * 128 + 4 * q);
                hacc[j] = fmaf(a[4 * q + 3], w.w, fmaf(a[4 * q + 2], w.z, fmaf(a[4 * q + 1], w.y, fmaf(a[4 * q], w.x, hacc[j]))));
              }
            }
          } else if (g.out_mode == TC_OUT_F32) {
            uint8_t *slab = slabs + (slab_i++ % TC_SLABS) * TC_SLAB_BYTES;
            if (lane == 0) bulk_wait_read<TC_SLABS - 1>();    // the store that last used this slab has read it
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 4; ++q)
              *reinterpret_cast<float4 *>(slab + lane * 64 + 16 * q) = make_float4(a[4 * q], a[4 * q + 1], a[4 * q + 2], a[4 * q + 3]);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) { if (g.debug != 1) tma_store_2d(&tmC_hi, slab, ncol, rbase); bulk_commit(); }
          } else {
            uint32_t hi[8], lo[8];
            if (g.fp16) {
#pragma unroll
              for (int q = 0; q < 8; ++q) split2_f16(a[2 * q], a[2 * q + 1], hi[q], lo[q]);
            } else {
#pragma unroll
              for (int q = 0; q < 8; ++q) split2_bf16(a[2 * q], a[2 * q + 1], hi[q], lo[q]);
            }
            const bool want_lo = g.c_plane != 0;          // c_plane == 0: the consumer only reads the hi plane
            uint8_t *slab = slabs + (slab_i++ % TC_SLABS) * TC_SLAB_BYTES;
            if (lane == 0) bulk_wait_read<TC_SLABS - 1>();    // the store that last used this slab has read it
            __syncwarp();
            *reinterpret_cast<uint4 *>(slab + lane * 32) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4 *>(slab + lane * 32 + 16) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
            if (want_lo) {
              *reinterpret_cast<uint4 *>(slab + 1024 + lane * 32) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
              *reinterpret_cast<uint4 *>(slab + 1024 + lane * 32 + 16) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              if (g.debug != 1) {
                tma_store_2d(&tmC_hi, slab, ncol, rbase);
                if (want_lo) tma_store_2d(&tmC_lo, slab + 1024, ncol, rbase);
              }
              bulk_commit();
            }
          }
        }
        if (BN == 128 && g.out_mode == TC_OUT_HEAD) {
          const int rt = wq * 32 + lane;                  // row inside the tile
          if (half == 1) *reinterpret_cast<float4 *>(sX + rt * 4) = make_float4(hacc[0], hacc[1], hacc[2], hacc[3]);
          named_bar_sync(3, 32 * TC_EPI_WARPS);
          if (half == 0 && row < g.M && g.debug != 1) {
            const float4 o = *reinterpret_cast<const float4 *>(sX + rt * 4);
            const float4 bb = *reinterpret_cast<const float4 *>(sW + 2 * 4 * 128 + (second ? 4 : 0));
            const float y0 = hacc[0] + o.x + bb.x, y1 = hacc[1] + o.y + bb.y, y2 = hacc[2] + o.z + bb.z, y3 = hacc[3] + o.w + bb.w;
            if (z == 0) *reinterpret_cast<float4 *>(g.pred_r + (size_t)row * 4) = make_float4(y0, y1, y2, y3);
            else if (z == 1) { float *d = g.pred_t + (size_t)row * 3; d[0] = y0; d[1] = y1; d[2] = y2; }
            else g.pred_c[row] = 1.0f / (1.0f + expf(-y0));
          }
        }
      } else {
#pragma unroll 1
      for (int c0 = half * TC_EPI_COLS; c0 < BN; c0 += 2 * TC_EPI_COLS) {
        if (n0 + c0 >= g.N || g.debug == 2) break;        // warp-uniform
        const int nc = n0 + c0;
        // TMEM -> registers (thread = row) -> smem slab [32][20] -> registers (lanes along columns),
        // so that every global store covers whole 32-byte sectors of a row.
        uint32_t raw[16];
        tmem_ld_32x16(tmem_base + (static_cast<uint32_t>(wq * 32) << 16) + as * BN + c0, raw);
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4 *>(sT + lane * TC_EPI_PITCH + 4 * q) = make_uint4(raw[4 * q], raw[4 * q + 1], raw[4 * q + 2], raw[4 * q + 3]);
        __syncwarp();
        if (g.out_mode != TC_OUT_BF16X2) {
          const int cq = lane & 3, rq = lane >> 2;        // 4 column quads x 8 rows per pass
          const int n = nc + 4 * cq;
          const float4 b0 = *reinterpret_cast<const float4 *>(sB + c0 + 4 * cq);
          const float4 b1 = *reinterpret_cast<const float4 *>(sB + BN + c0 + 4 * cq);
          float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int rr = 8 * i + rq, row = rbase + rr;
            const float4 a4 = *reinterpret_cast<const float4 *>(sT + rr * TC_EPI_PITCH + 4 * cq);
            const bool second = row >= split_row;
            const float4 bb = second ? b1 : b0;
            float a[4] = {a4.x + bb.x, a4.y + bb.y, a4.z + bb.z, a4.w + bb.w};
            if (g.relu) {
#pragma unroll
              for (int j = 0; j < 4; ++j) a[j] = fmaxf(a[j], 0.f);
            }
            if (row < g.M && g.debug != 1) {
              if (g.out_mode == TC_OUT_F32) {
                float *dst = g.Cf + (long long)row * g.ldc + z * g.c_col_z + n;
                if (n + 3 < g.N && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                  *reinterpret_cast<float4 *>(dst) = make_float4(a[0], a[1], a[2], a[3]);
                } else {
#pragma unroll
                  for (int j = 0; j < 4; ++j) if (n + j < g.N) dst[j] = a[j];
                }
              } else {                                    // TC_OUT_POOL: <= 2 frames per tile
                if (!second) { sa[0] += a[0]; sa[1] += a[1]; sa[2] += a[2]; sa[3] += a[3]; }
                else         { sb[0] += a[0]; sb[1] += a[1]; sb[2] += a[2]; sb[3] += a[3]; }
              }
            }
          }
          if (g.out_mode == TC_OUT_POOL && rbase < g.M) {
            // pool[(row group)][slot] : slot 0 = frame of the group's first row, slot 1 = the next frame
            const bool first_is_a = rbase < split_row;    // warp-uniform
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#pragma unroll
              for (int o = 4; o < 32; o <<= 1) {
                sa[j] += __shfl_xor_sync(0xffffffffu, sa[j], o);
                sb[j] += __shfl_xor_sync(0xffffffffu, sb[j], o);
              }
            }
            if (rq == 0) {
              float *dst = g.pool + ((long long)(rbase >> 5) * 2) * g.N + n;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (n + j < g.N) {
                  dst[j] = first_is_a ? sa[j] : sb[j];
                  dst[g.N + j] = first_is_a ? sb[j] : 0.f;
                }
            }
          }
        } else {
          const int cq = lane & 1, rq = lane >> 1;        // 2 column octets x 16 rows per pass
          const int n = nc + 8 * cq;
          float b0[8], b1[8];
          *reinterpret_cast<float4 *>(b0) = *reinterpret_cast<const float4 *>(sB + c0 + 8 * cq);
          *reinterpret_cast<float4 *>(b0 + 4) = *reinterpret_cast<const float4 *>(sB + c0 + 8 * cq + 4);
          *reinterpret_cast<float4 *>(b1) = *reinterpret_cast<const float4 *>(sB + BN + c0 + 8 * cq);
          *reinterpret_cast<float4 *>(b1 + 4) = *reinterpret_cast<const float4 *>(sB + BN + c0 + 8 * cq + 4);
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int rr = 16 * i + rq, row = rbase + rr;
            const float4 a0 = *reinterpret_cast<const float4 *>(sT + rr * TC_EPI_PITCH + 8 * cq);
            const float4 a1 = *reinterpret_cast<const float4 *>(sT + rr * TC_EPI_PITCH + 8 * cq + 4);
            float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const bool second = row >= split_row;
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              a[j] += second ? b1[j] : b0[j];
              if (g.relu) a[j] = fmaxf(a[j], 0.f);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (g.fp16) split2_f16(a[2 * q], a[2 * q + 1], hi[q], lo[q]);
              else split2_bf16(a[2 * q], a[2 * q + 1], hi[q], lo[q]);
            }
            if (row < g.M && g.debug != 1) {
              __nv_bfloat16 *dh = g.Cb + (long long)row * g.ldcb + z * g.c_col_z + n;
              __nv_bfloat16 *dl = dh + g.c_plane;
              const bool want_lo = g.c_plane != 0;        // c_plane == 0: the consumer only reads the hi plane
              if (n + 7 < g.N) {
                *reinterpret_cast<uint4 *>(dh) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                if (want_lo) *reinterpret_cast<uint4 *>(dl) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (n + j < g.N) {
                    dh[j] = __ushort_as_bfloat16((unsigned short)(hi[j >> 1] >> ((j & 1) * 16)));
                    if (want_lo) dl[j] = __ushort_as_bfloat16((unsigned short)(lo[j >> 1] >> ((j & 1) * 16)));
                  }
              }
            }
          }
        }
        __syncwarp();
      }
      }
      tmem_ld_wait();                                     // a prefetched chunk that was never consumed (no valid columns)
      fence_before_sync();
      __syncwarp();
      if (lane == 0) {                                    // one arrival per epilogue warp (CG * 8 in total)
        if (CG == 2) mbar_arrive_cluster(mapa_u32(smem_u32(tempty_bar + as), 0));
        else mbar_arrive(tempty_bar + as);
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }

  if (warp >= 4 && lane == 0) bulk_wait_all();            // outstanding TMA stores of this thread
  fence_before_sync();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    fence_after_sync();
    if (CG == 2) tmem_dealloc_cg2(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------------------------
// ap[f][c] = (1/N) * sum of the 32-row-group partials of frame f.  Block = 64 float4 columns x 4 group
// lanes (each lane takes every 4th group, 16-byte loads, then a fixed-order combine).  Slot 1 of a group
// is only ever read for the group that straddles the start of frame f (the GEMM epilogue writes it only then).
__global__ void __launch_bounds__(256)
pool_finish_kernel(const float *__restrict__ pool, int rows_per_frame, int M, int ncols, float *__restrict__ ap,
                   __nv_bfloat16 *__restrict__ apb, long long apb_plane, int fp16) {
  __shared__ float4 red[4][64];
  const int f = blockIdx.y;
  const int cl = threadIdx.x & 63, gl = threadIdx.x >> 6;
  const int c = (blockIdx.x * 64 + cl) * 4;
  const long long r_beg = (long long)f * rows_per_frame;
  long long r_end = r_beg + rows_per_frame;
  if (r_end > M) r_end = M;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c < ncols) {
    const long long g0 = r_beg >> 5, g1 = (r_end - 1) >> 5;
#pragma unroll 4
    for (long long g = g0 + gl; g <= g1; g += 4) {
      const int fa = (int)((g << 5) / rows_per_frame);        // frame of the group's first row
      const float4 v = __ldg(reinterpret_cast<const float4 *>(pool + (g * 2 + (fa == f ? 0 : 1)) * ncols + c));
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  red[gl][cl] = s;
  __syncthreads();
  if (gl == 0 && c < ncols) {
    float4 t = red[0][cl];
#pragma unroll
    for (int i = 1; i < 4; ++i) { t.x += red[i][cl].x; t.y += red[i][cl].y; t.z += red[i][cl].z; t.w += red[i][cl].w; }
    const float inv = 1.0f / (float)rows_per_frame;
    const float4 o = make_float4(t.x * inv, t.y * inv, t.z * inv, t.w * inv);
    *reinterpret_cast<float4 *>(ap + (long long)f * ncols + c) = o;
    if (apb) {                                                  // operand planes of the M = frames GEMMs
      uint2 ph, pl;
      if (fp16) { split2_f16(o.x, o.y, ph.x, pl.x); split2_f16(o.z, o.w, ph.y, pl.y); }
      else { split2_bf16(o.x, o.y, ph.x, pl.x); split2_bf16(o.z, o.w, ph.y, pl.y); }
      *reinterpret_cast<uint2 *>(apb + (long long)f * ncols + c) = ph;
      *reinterpret_cast<uint2 *>(apb + apb_plane + (long long)f * ncols + c) = pl;
    }
  }
}

int pool_finish(const float *pool, int frames, int rows_per_frame, int M, int ncols, float *ap,
                __nv_bfloat16 *apb, long long apb_plane, int fp16, cudaStream_t st) {
  if (ncols % 4) { set_error("pool_finish: ncols must be a multiple of 4"); return PLR2_E_SHAPE; }
  dim3 grid((ncols / 4 + 63) / 64, frames);
  pool_finish_kernel<<<grid, 256, 0, st>>>(pool, rows_per_frame, M, ncols, ap, apb, apb_plane, fp16);
  return check_launch("pool_finish");
}

__global__ void __launch_bounds__(256)
split_bf16x2_kernel(const float *__restrict__ src, long long ld, long long rows, int cols,
                    __nv_bfloat16 *__restrict__ dst, long long ldd, long long plane) {
  const long long total = rows * cols;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols; const int c = (int)(i % cols);
    const float v = src[r * ld + c];
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    dst[r * ldd + c] = h;
    dst[plane + r * ldd + c] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

int split_bf16x2(const float *src, long long ld, long long rows, int cols, __nv_bfloat16 *dst, long long ldd,
                 long long plane, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  long long blocks = (rows * cols + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  split_bf16x2_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, ld, rows, cols, dst, ldd, plane);
  return check_launch("split_bf16x2");
}

// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// K-major bf16 matrix [rows][cols] with row pitch ld (elements); box = 64 columns x box_rows rows.
static int make_map(CUtensorMap *m, const __nv_bfloat16 *base, long long cols, long long rows, long long ld,
                    int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled unavailable"); return PLR2_E_UNSUPPORTED; }
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld * 2) % 16) { set_error("tensor map: base/pitch not 16-byte aligned"); return PLR2_E_ALIGN; }
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16 *>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d): cols=%lld rows=%lld ld=%lld", (int)r, cols, rows, ld); return PLR2_E_UNSUPPORTED; }
  return 0;
}

// MN-major operand: bf16 matrix [k_rows][mn] with row pitch ld; box = 64 mn-columns x 64 k-rows (one 8 KB block of
// the shared-memory tile, make_desc_mn128).  Columns >= mn and rows >= k_rows read as zero.
static int make_map_mn(CUtensorMap *m, const __nv_bfloat16 *base, long long mn, long long k_rows, long long ld) {
  return make_map(m, base, mn, k_rows, ld, TC_BK);
}

// Row-major output matrix [rows][cols] (pitch ld elements) written in boxes of 16 columns x 32 rows from a
// dense (unswizzled) shared-memory slab.
static int make_store_map(CUtensorMap *m, void *base, bool f32, long long cols, long long rows, long long ld) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled unavailable"); return PLR2_E_UNSUPPORTED; }
  const int esz = f32 ? 4 : 2;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)TC_EPI_COLS, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstride, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (store) failed (%d): cols=%lld rows=%lld ld=%lld", (int)r, cols, rows, ld); return PLR2_E_UNSUPPORTED; }
  return 0;
}

static int sm_count() { return device_sm_count(); }

// Optional CUDA-event timing of every tensor-core launch (plr2_profile_begin / plr2_profile_end).
struct TcProfile {
  bool on = false;
  std::vector<cudaEvent_t> ev;      // pairs
  size_t used = 0;
  double flops = 0.0;
  double executed = 0.0;   // tensor-core FLOPs actually issued (2 or 3 MMAs per algorithmic MAC)
  long long launches = 0;
};
static TcProfile g_prof;

static cudaEvent_t prof_event() {
  if (g_prof.used == g_prof.ev.size()) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    g_prof.ev.push_back(e);
  }
  return g_prof.ev[g_prof.used++];
}

static int max_pairs_resident(const void *fn, int smem_bytes) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * sm_count(), 1, 1);
  cfg.blockDim = dim3(TC_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, fn, &cfg) != cudaSuccess || n <= 0) { cudaGetLastError(); n = sm_count() / 2; }
  return n;
}

template <int BN, int NPROD, int CG>
static int launch_tc(const TcGemm &g, cudaStream_t st) {
  using Cfg = TcCfg<BN, NPROD, CG>;
  CUtensorMap ma_hi, ma_lo, mb_hi, mb_lo;
  const long long a_cols = (long long)g.a_col_z * (g.batch - 1) + g.K;
  const long long w_rows = (long long)g.b_row_z * (g.batch - 1) + g.N;
  int rc;
  if (g.mn_major) {
    const long long k_rows = g.k_rows > 0 ? g.k_rows : a_cols;     // a_cols: k extent over all problems z
    if ((rc = make_map_mn(&ma_hi, g.A, g.M, k_rows, g.lda))) return rc;
    if ((rc = make_map_mn(&ma_lo, g.A + (NPROD == 3 ? g.a_plane : 0), g.M, k_rows, g.lda))) return rc;
    if ((rc = make_map_mn(&mb_hi, g.W, w_rows, k_rows, g.ldw))) return rc;
    if ((rc = make_map_mn(&mb_lo, g.W + (NPROD == 1 ? 0 : g.w_plane), w_rows, k_rows, g.ldw))) return rc;
  } else {
  if ((rc = make_map(&ma_hi, g.A, a_cols, g.M, g.lda, TC_BM))) return rc;
  if ((rc = make_map(&ma_lo, g.A + (NPROD == 3 ? g.a_plane : 0), a_cols, g.M, g.lda, TC_BM))) return rc;
  if ((rc = make_map(&mb_hi, g.W, g.K, w_rows, g.ldw, Cfg::B_ROWS))) return rc;
  if ((rc = make_map(&mb_lo, g.W + (NPROD == 1 ? 0 : g.w_plane), g.K, w_rows, g.ldw, Cfg::B_ROWS))) return rc;
  }
  // Output maps: TMA stores need 16-byte aligned base and pitch; with batch > 1 a partial last chunk would
  // spill into the next problem's columns, so N must then be a multiple of the chunk width.
  CUtensorMap mc_hi = ma_hi, mc_lo = ma_hi;
  int tma_store = 0;
  const long long c_cols = g.c_col_z * (g.batch - 1) + g.N;
  if (g.out_mode == TC_OUT_F32) {
    if ((g.ldc % 4) == 0 && (reinterpret_cast<uintptr_t>(g.Cf) & 15) == 0 && (g.batch == 1 || g.N % TC_EPI_COLS == 0)) {
      if ((rc = make_store_map(&mc_hi, g.Cf, true, c_cols, g.M, g.ldc))) return rc;
      tma_store = 1;
    }
  } else if (g.out_mode == TC_OUT_BF16X2) {
    if ((reinterpret_cast<uintptr_t>(g.Cb) & 15) == 0 && (g.c_plane % 8) == 0 && (g.batch == 1 || g.N % TC_EPI_COLS == 0)) {
      if ((rc = make_store_map(&mc_hi, g.Cb, false, c_cols, g.M, g.ldcb))) return rc;
      if (g.c_plane != 0 && (rc = make_store_map(&mc_lo, g.Cb + g.c_plane, false, c_cols, g.M, g.ldcb))) return rc;
      tma_store = 1;
    }
  }
  TcKernelArgs k{};
  k.tma_store = tma_store;
  k.fp16 = g.fp16; k.mn_major = g.mn_major;
  k.M = g.M; k.N = g.N; k.K = g.K; k.batch = g.batch;
  k.m_tiles = (g.M + TC_BM * CG - 1) / (TC_BM * CG); k.n_tiles = (g.N + BN - 1) / BN;
  k.a_col_z = g.a_col_z; k.b_row_z = g.b_row_z; k.c_col_z = g.c_col_z; k.bias_z = g.bias_z;
  k.bias = g.bias; k.frame_bias = g.frame_bias; k.rows_per_frame = g.rows_per_frame > 0 ? g.rows_per_frame : 1; k.ldfb = g.ldfb;
  k.relu = g.relu; k.out_mode = g.out_mode;
  k.Cf = g.Cf; k.ldc = g.ldc; k.Cb = g.Cb; k.ldcb = g.ldcb; k.c_plane = g.c_plane; k.pool = g.pool;
  k.w4 = g.w4; k.b4 = g.b4; k.obj = g.obj; k.num_obj = g.num_obj; k.pred_r = g.pred_r; k.pred_t = g.pred_t; k.pred_c = g.pred_c;
  { static int dbg = -1; if (dbg < 0) { const char *e = getenv("PLR2_TC_DEBUG"); dbg = e ? atoi(e) : 0; } k.debug = dbg; }
  auto kfn = gemm_bf16x2_kernel<BN, NPROD, CG>;
  // per instantiation AND device: CTAs (CG == 1) or CTA pairs (CG == 2) resident at once; the shared-memory
  // attribute is a per-device property of the function
  static PerDeviceInt units_cache;
  const int dev = current_device();
  int units_max = units_cache.get(dev);
  if (!units_max) {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) { set_error("gemm_bf16x2: smem attribute: %s", cudaGetErrorString(e)); return (int)e; }
    units_max = CG == 2 ? max_pairs_resident((const void *)kfn, Cfg::SMEM_BYTES) : sm_count();
    units_cache.set(dev, units_max);
  }
  const long long tiles = (long long)k.m_tiles * k.n_tiles * k.batch;
  const int units = (int)(tiles < units_max ? tiles : units_max);
  if (g_prof.on) cudaEventRecord(prof_event(), st);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(units * CG, 1, 1);
  cfg.blockDim = dim3(TC_THREADS, 1, 1);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kfn, ma_hi, ma_lo, mb_hi, mb_lo, mc_hi, mc_lo, k);
  if (e != cudaSuccess) { set_error("gemm_bf16x2: launch: %s", cudaGetErrorString(e)); return (int)e; }
  if (g_prof.on) {
    cudaEventRecord(prof_event(), st);
    g_prof.flops += 2.0 * g.M * g.N * g.K * g.batch;
    g_prof.executed += 2.0 * g.M * g.N * g.K * g.batch * NPROD;
    g_prof.launches += 1;
  }
  return check_launch("gemm_bf16x2");
}

static int tc_cta_group() {             // PLR2_TC_CTA_GROUP=1 selects the single-CTA kernel (debugging aid)
  static int cg = 0;
  if (!cg) {
    const char *e = getenv("PLR2_TC_CTA_GROUP");
    cg = (e && e[0] == '1') ? 1 : 2;
  }
  return cg;
}

int gemm_bf16x2(const TcGemm &g, cudaStream_t st) {
  if (g.M <= 0 || g.N <= 0) return 0;
  if (g.K <= 0 || g.K % TC_BK) { set_error("gemm_bf16x2: K=%d must be a positive multiple of %d", g.K, TC_BK); return PLR2_E_SHAPE; }
  if (!g.A || !g.W) { set_error("gemm_bf16x2: null operand"); return PLR2_E_NULLPTR; }
  if (g.out_mode == TC_OUT_F32 && !g.Cf) { set_error("gemm_bf16x2: null Cf"); return PLR2_E_NULLPTR; }
  if (g.out_mode == TC_OUT_BF16X2 && (!g.Cb || (g.ldcb % 8) || (g.c_col_z % 8))) { set_error("gemm_bf16x2: bf16 output needs 16-byte aligned rows"); return PLR2_E_ALIGN; }
  if (g.out_mode == TC_OUT_POOL && (!g.pool || g.batch != 1)) { set_error("gemm_bf16x2: pooled output needs batch == 1"); return PLR2_E_UNSUPPORTED; }
  if (g.out_mode == TC_OUT_HEAD && (g.N != 128 || g.batch != 3 || !g.w4 || !g.b4 || !g.obj || !g.pred_r || !g.pred_t || !g.pred_c ||
                                    g.num_obj <= 0 || (reinterpret_cast<uintptr_t>(g.pred_r) & 15))) {
    set_error("gemm_bf16x2: fused head output needs N == 128, batch == 3 (r|t|c) and all head pointers");
    return PLR2_E_UNSUPPORTED;
  }
  if ((g.out_mode == TC_OUT_POOL || g.out_mode == TC_OUT_HEAD || g.frame_bias) && g.rows_per_frame < TC_BM) { set_error("gemm_bf16x2: per-frame epilogues need rows_per_frame >= %d", TC_BM); return PLR2_E_UNSUPPORTED; }
  if (g.mn_major && (g.out_mode != TC_OUT_F32 || g.a_hi_only || g.frame_bias)) { set_error("gemm_bf16x2: MN-major operands are for plain FP32-output products"); return PLR2_E_UNSUPPORTED; }
  const bool hi_only = g.a_hi_only != 0;
  if (g.w_hi_only && !hi_only) { set_error("gemm_bf16x2: w_hi_only needs a_hi_only (single product)"); return PLR2_E_UNSUPPORTED; }
  if (g.w_hi_only) {              // single product X_hi W_hi (large layers only: 256-wide tiles on CTA pairs)
    if (tc_cta_group() == 2) return g.N >= 256 ? launch_tc<256, 1, 2>(g, st) : launch_tc<128, 1, 2>(g, st);
    return g.N >= 256 ? launch_tc<256, 1, 1>(g, st) : launch_tc<128, 1, 1>(g, st);
  }
  if (tc_cta_group() == 2) {
    // few 256-wide tiles (M = frames problems): halve the tile so twice as many CTA pairs share the work
    const int pairs = sm_count() / 2;
    const long long tiles256 = (long long)((g.M + 2 * TC_BM - 1) / (2 * TC_BM)) * ((g.N + 255) / 256) * g.batch;
    static int allow_narrow = -1;
    if (allow_narrow < 0) { const char *e = getenv("PLR2_TC_NARROW"); allow_narrow = (e && e[0] == '0') ? 0 : 1; }
    const bool narrow = g.N < 256 || (allow_narrow && 2 * tiles256 <= pairs && g.out_mode != TC_OUT_POOL);
    if (!narrow) return hi_only ? launch_tc<256, 2, 2>(g, st) : launch_tc<256, 3, 2>(g, st);
    return hi_only ? launch_tc<128, 2, 2>(g, st) : launch_tc<128, 3, 2>(g, st);
  }
  if (g.N >= 256) return hi_only ? launch_tc<256, 2, 1>(g, st) : launch_tc<256, 3, 1>(g, st);
  return hi_only ? launch_tc<128, 2, 1>(g, st) : launch_tc<128, 3, 1>(g, st);
}

}  // namespace plr2

// C-ABI test/primitive entries ------------------------------------------------------------------
extern "C" int plr2_split_bf16x2(const float *src, long long ld, long long rows, int cols, void *dst_planes,
                                 long long ldd, long long plane_stride, void *stream) {
  if (!src || !dst_planes) { plr2::set_error("plr2_split_bf16x2: null pointer"); return PLR2_E_NULLPTR; }
  return plr2::split_bf16x2(src, ld, rows, cols, (__nv_bfloat16 *)dst_planes, ldd, plane_stride, (cudaStream_t)stream);
}

extern "C" int plr2_gemm_bf16x2(const void *A_planes, long long lda, long long a_plane, const void *W_planes,
                                long long ldw, long long w_plane, int M, int N, int K, const float *bias, int relu,
                                float *C, long long ldc, void *stream) {
  plr2::TcGemm g;
  g.A = (const __nv_bfloat16 *)A_planes; g.lda = lda; g.a_plane = a_plane;
  g.W = (const __nv_bfloat16 *)W_planes; g.ldw = ldw; g.w_plane = w_plane;
  g.M = M; g.N = N; g.K = K; g.bias = bias; g.relu = relu; g.out_mode = plr2::TC_OUT_F32; g.Cf = C; g.ldc = ldc;
  return plr2::gemm_bf16x2(g, (cudaStream_t)stream);
}

// Weight-gradient form on MN-major planes: C[M][N] (ldc) = A^T B with A = planes [k_rows][lda] (M columns used),
// B = planes [k_rows][ldb] (N columns used); k_rows need not be a multiple of 64 (the tail reads as zero).
extern "C" int plr2_gemm_bf16x2_tn(const void *A_planes, long long lda, long long a_plane, const void *B_planes,
                                   long long ldb, long long b_plane, int M, int N, int k_rows, float *C,
                                   long long ldc, void *stream) {
  if (k_rows <= 0) { plr2::set_error("plr2_gemm_bf16x2_tn: k_rows=%d", k_rows); return PLR2_E_SHAPE; }
  plr2::TcGemm g;
  g.A = (const __nv_bfloat16 *)A_planes; g.lda = lda; g.a_plane = a_plane;
  g.W = (const __nv_bfloat16 *)B_planes; g.ldw = ldb; g.w_plane = b_plane;
  g.M = M; g.N = N; g.K = (k_rows + 63) / 64 * 64; g.k_rows = k_rows; g.mn_major = 1;
  g.out_mode = plr2::TC_OUT_F32; g.Cf = C; g.ldc = ldc;
  return plr2::gemm_bf16x2(g, (cudaStream_t)stream);
}

// Profiling of the tensor-core launches: between begin and end every gemm_bf16x2 launch is bracketed
// by CUDA events on its stream.  end() synchronises the device and returns the summed kernel time,
// the number of launches and their algorithmic FLOPs (2 M N K per problem; the kernel executes 3x
// that many bf16 tensor-core FLOPs).
extern "C" int plr2_profile_begin(void) {
  plr2::g_prof.on = true; plr2::g_prof.used = 0; plr2::g_prof.flops = 0.0; plr2::g_prof.executed = 0.0; plr2::g_prof.launches = 0;
  return 0;
}
extern "C" int plr2_profile_end(double *ms, long long *launches, double *flops) {
  using namespace plr2;
  g_prof.on = false;
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { set_error("plr2_profile_end: %s", cudaGetErrorString(e)); return (int)e; }
  double total = 0.0;
  for (size_t i = 0; i + 1 < g_prof.used; i += 2) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, g_prof.ev[i], g_prof.ev[i + 1]) == cudaSuccess) total += t;
  }
  if (ms) *ms = total;
  if (launches) *launches = g_prof.launches;
  if (flops) *flops = g_prof.flops;
  return 0;
}
// bf16 tensor-core FLOPs issued by the launches of the last begin/end window (>= the algorithmic FLOPs:
// split-bf16 layers cost 3 MMAs per MAC, hi-only-activation layers 2).
extern "C" double plr2_profile_executed_flops(void) { return plr2::g_prof.executed; }
