// Internal GEMM interface shared by gemm_f32.cu (FP32 SIMT), gemm_tf32.cu (tcgen05) and the network
// orchestration (posenet.cu).
#pragma once
#include "common.cuh"

namespace plr2 {

// C(m,n) = epi( [accumulate ? C_old : 0] + sum_k A(m,k) B(k,n) + bias[n] + frame_bias[m/rpf][n] )
// epi: optional ReLU, then optional multiply by (relu_mask(m,n) > 0).
// `batch` > 1 runs independent problems z = 0..batch-1 with all pointers advanced by the z* strides
// (used for the r/t/c heads).
struct GemmArgs {
  const float *A = nullptr; long long sam = 0, sak = 0;
  const float *B = nullptr; long long sbk = 0, sbn = 0;
  float *C = nullptr; long long ldc = 0;
  int M = 0, N = 0, K = 0;
  const float *bias = nullptr;
  const float *frame_bias = nullptr; int rows_per_frame = 1; long long ldfb = 0;
  const float *relu_mask = nullptr; long long ldmask = 0;
  int relu = 0, accumulate = 0;
  int batch = 1; long long zA = 0, zB = 0, zC = 0, zbias = 0, zfb = 0, zmask = 0;
  int k_per_split = 0, splits = 1;   // filled by the launcher
  // Tensor-core training path only: A's split planes may be taken from the scratch if the immediately preceding
  // weight-gradient call on this stream packed the SAME matrix (dW = dY^T X then dX = dY W: dY is split once).
  // A hint - honoured only when the scratch record matches pointer, shape and pitch; the caller guarantees that
  // the matrix was not written in between.
  int reuse_a_planes = 0;
  // Weight gradients on the tensor-core path: column sums of A (= dY: the bias gradient) per 32-row tile, written by
  // the pass that splits dY - colsum_part [tiles][M * batch] (capacity colsum_cap floats); *colsum_tiles is set to the
  // number of tiles when the partials were written (left untouched otherwise: the caller then sums dY itself).
  float *colsum_part = nullptr; long long colsum_cap = 0; int *colsum_tiles = nullptr;
};

// FP32 SIMT GEMM.  split_k > 1 adds partial sums atomically into C (needs accumulate=1, no relu/mask).
int gemm_f32(GemmArgs g, int split_k, cudaStream_t st);

// convenience builders ------------------------------------------------------------------------
// Y[M, N] (ldc) = X[M, K] (lda, k contiguous) * W[N, K]^T (ldw, k contiguous)      "forward"
inline GemmArgs gemm_nt(const float *X, long long lda, const float *W, long long ldw, float *Y,
                        long long ldc, int M, int N, int K) {
  GemmArgs g;
  g.A = X; g.sam = lda; g.sak = 1;
  g.B = W; g.sbk = 1; g.sbn = ldw;
  g.C = Y; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  return g;
}
// dX[M, N] = dY[M, K] (lda) * W[K, N] (ldw, n contiguous)                           "input grad"
inline GemmArgs gemm_nn(const float *dY, long long lda, const float *W, long long ldw, float *dX,
                        long long ldc, int M, int N, int K) {
  GemmArgs g;
  g.A = dY; g.sam = lda; g.sak = 1;
  g.B = W; g.sbk = ldw; g.sbn = 1;
  g.C = dX; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  return g;
}
// dW[M, N] (ldc) += dY[K, M]^T (lda, m contiguous) * X[K, N] (ldb, n contiguous)    "weight grad"
inline GemmArgs gemm_tn(const float *dY, long long lda, const float *X, long long ldb, float *dW,
                        long long ldc, int M, int N, int K) {
  GemmArgs g;
  g.A = dY; g.sam = 1; g.sak = lda;
  g.B = X; g.sbk = ldb; g.sbn = 1;
  g.C = dW; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  g.accumulate = 1;
  return g;
}

}  // namespace plr2
