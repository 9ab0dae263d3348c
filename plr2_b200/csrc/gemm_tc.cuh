// Internal interface of the tcgen05 split-bf16 GEMM (gemm_bf16x2.cu).
#pragma once
#include "common.cuh"
#include <cuda_bf16.h>

namespace plr2 {

enum { TC_OUT_F32 = 0, TC_OUT_BF16X2 = 1, TC_OUT_POOL = 2, TC_OUT_HEAD = 3 };

// Y = epi(X W^T).  X: bf16 planes [2][rows][lda] (plane stride a_plane elements), W likewise.
// batch > 1: problem z reads X columns [z*a_col_z, +K), W rows [z*b_row_z, +N), writes C columns
// [z*c_col_z, +N) and uses bias + z*bias_z.
struct TcGemm {
  const __nv_bfloat16 *A = nullptr; long long lda = 0, a_plane = 0;
  const __nv_bfloat16 *W = nullptr; long long ldw = 0, w_plane = 0;
  int M = 0, N = 0, K = 0, batch = 1;
  int a_col_z = 0, b_row_z = 0; long long c_col_z = 0, bias_z = 0;
  const float *bias = nullptr, *frame_bias = nullptr; int rows_per_frame = 1; long long ldfb = 0;
  int relu = 0, out_mode = TC_OUT_F32;
  int a_hi_only = 0;          // 1: use only the hi plane of X (2 MMAs per MAC; for layers whose output is only pooled)
  int w_hi_only = 0;          // 1 (with a_hi_only): single product X_hi W_hi (1 MMA per MAC)
  // 1: both operands are stored MN-major - A as [k rows][M columns] (pitch lda), W as [k rows][N columns] (pitch
  // ldw): C = A^T W, the weight-gradient form.  Problem z reads k-rows [z*a_col_z, +K) of BOTH operands and W
  // columns [z*b_row_z, +N); k_rows = number of valid k-rows in memory (rows beyond it read as zero), 0 = all.
  int mn_major = 0; long long k_rows = 0;
  int fp16 = 0;               // operand planes and TC_OUT_BF16X2 outputs are IEEE fp16 (11-bit mantissa) instead of bf16; same layout
  float *Cf = nullptr; long long ldc = 0;
  __nv_bfloat16 *Cb = nullptr; long long ldcb = 0, c_plane = 0;
  float *pool = nullptr;      // TC_OUT_POOL: [ceil(M/32)][2][N] partial column sums
  // TC_OUT_HEAD (N == 128, batch z = head r|t|c): the 128-wide ReLU output of head z is not stored; its last
  // layer is applied in the epilogue for the class of the row's frame only (conv4_* + select + sigmoid,
  // lib/network.py:117-130): w4 [(4+3+1)*num_obj][128] rows r|t|c, b4 likewise.
  const float *w4 = nullptr, *b4 = nullptr; const int64_t *obj = nullptr; int num_obj = 0;
  float *pred_r = nullptr, *pred_t = nullptr, *pred_c = nullptr;
};

int gemm_bf16x2(const TcGemm &g, cudaStream_t st);
// ap[f][c] = (1/rows_per_frame) * sum of the pooling partials belonging to frame f; apb (optional): the same
// values as split-bf16 planes [2][frames][ncols], plane stride apb_plane elements
int pool_finish(const float *pool, int frames, int rows_per_frame, int M, int ncols, float *ap,
                __nv_bfloat16 *apb, long long apb_plane, int fp16, cudaStream_t st);
// dst planes <- split(src): hi = bf16(v), lo = bf16(v - hi)
int split_bf16x2(const float *src, long long ld, long long rows, int cols, __nv_bfloat16 *dst, long long ldd,
                 long long plane, cudaStream_t st);


// Training path (gemm_tc_f32.cu): a GemmArgs problem (gemm.cuh) on FP32 operands run on the tcgen05 kernel
// through split-bf16 scratch planes.  pa / pb: two planes of `elems` bf16 each; ct: `elems` floats.
struct TcScratch { __nv_bfloat16 *pa = nullptr, *pb = nullptr; float *ct = nullptr; size_t elems = 0; };
struct GemmArgs;
bool gemm_tc_eligible(const GemmArgs &g);
int gemm_tc_f32(const GemmArgs &g, const TcScratch &S, cudaStream_t st);

}  // namespace plr2
