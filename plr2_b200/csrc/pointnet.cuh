// Internal declarations of the non-GEMM PointNet kernels (pointnet.cu).
#pragma once
#include "net_layout.cuh"
#include <cuda_bf16.h>

namespace plr2 {

int first_layer(const float *packed, const NetLayout &L, const float *out_img, const int64_t *choose,
                const float *emb_in, const float *x, int frames, int N, int hw, float *pf, float *emb_out,
                float *embT, const int64_t *obj, unsigned int *status, cudaStream_t st);
int first_layer_bf16(const float *packed, const NetLayout &L, const float *out_img, const int64_t *choose,
                     const float *emb_in, const float *x, int frames, int N, int hw, __nv_bfloat16 *pfb,
                     long long pf_plane, float *emb_out, int skip_e, const int64_t *obj, unsigned int *status,
                     cudaStream_t st);
int scatter_emb_grad(const float *demb, const int64_t *choose, int frames, int N, int hw, float *g_out_img,
                     cudaStream_t st);
int segment_colsum(const float *X, long long ld, int nseg, int rows_per_seg, int ncols, float scale, float *out,
                   long long ldo, cudaStream_t st);
// out[c] = sum over `rows` rows of X[r][c], one segment, 32 row lanes per block (partials of a bias gradient)
int colsum_tall(const float *X, long long ld, int rows, int ncols, float *out, cudaStream_t st);
int segment_colsum_ragged(const float *X, long long ld, int nseg, int rows_per_seg, size_t total_rows, int ncols, float *out,
                          long long ldo, cudaStream_t st);
int segment_bcast_mask(const float *dseg, const float *X, size_t rows, int rows_per_seg, int ncols, float scale,
                       float *dX, cudaStream_t st);
int select_outputs(const float *raw, const NetLayout &L, const int64_t *obj, int rows_per_frame, size_t rows,
                   float *pred_r, float *pred_t, float *pred_c, cudaStream_t st);
int select_outputs_bwd(float *raw, const NetLayout &L, const int64_t *obj, int rows_per_frame, size_t rows,
                       const float *g_r, const float *g_t, const float *g_c, cudaStream_t st);
// heads.cu
int frame_linear(const float *X, long long ldx, const float *W, long long ldw, const float *bias, float *Y,
                 long long ldy, int M, int N, int K, int relu, int batch, long long zX, long long zW, long long zB,
                 long long zY, cudaStream_t st);
int head_out(const float *H, long long ldh, const float *W4, const float *b4, const int64_t *obj, int num_obj,
             int frames, int rows_per_frame, float *pred_r, float *pred_t, float *pred_c, cudaStream_t st);
int net_copy_params(const NetLayout &L, const plr2_param_table *P, float *packed, bool to_packed, int round,
                    cudaStream_t st, int accumulate = 0);

}  // namespace plr2
