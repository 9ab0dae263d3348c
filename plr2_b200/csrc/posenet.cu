// Orchestration of kernels (1)+(2): PoseNet / PoseRefineNet forward and backward as a fixed sequence
// of launches on one stream (no host synchronisation, CUDA-graph capturable).
// Reference: lib/network.py:53-68,95-132 (PoseNet), :151-168,187-206 (PoseRefineNet).
//
// Activations are POINT-MAJOR: one row per point (row = frame*N + n), channels contiguous, so every
// Conv1d(k=1) is Y = X W^T with both operands K-contiguous.  Feature buffer pf [R,384]:
//   [0,64) relu(conv1 x) | [64,128) relu(e_conv1 emb) | [128,256) relu(conv2 .) | [256,384) relu(e_conv2 .)
// i.e. pointfeat_1 | pointfeat_2 of the reference, already concatenated.
// Algebraic hoist: the 1024 global-feature inputs of conv1_{r,t,c} are constant per
// frame, so they enter as a per-frame bias hb = W1[:,384:] ap + b1 instead of 1024 extra K per point.
#include "gemm.cuh"
#include "gemm_tc.cuh"
#include "pointnet.cuh"

namespace plr2 {

// Where a contraction of the FP32-activation path (training, or inference without the bf16 workspace) runs:
// on the tcgen05 kernel through split-bf16 scratch planes when the caller asked for PLR2_PREC_BF16X2 and the
// call keeps activations for backward (gemm_tc_f32.cu), else on the FP32 SIMT kernel.
struct GemmCtx {
  bool tc = false;
  TcScratch S;
};
static bool prec_is_tc(int precision) { return precision == PLR2_PREC_BF16X2 || precision == PLR2_PREC_BF16X2_ALL; }
// forward = the call is a training FORWARD.  Its ReLU outputs are the masks of the backward: an operand error
// of 2^-17 flips the sign of about one pre-activation in 10^5 against an FP32 evaluation, and every flip moves the
// gradients of that point by O(1) (measured: parameter gradients of a 500-point frame only 92-98 % within rtol
// 1e-3 instead of 100 %).  PLR2_PREC_BF16X2 therefore keeps the training forward on the FP32 kernel and runs the
// backward (no discontinuity: errors stay ~1e-5) on the tensor cores; PLR2_PREC_BF16X2_ALL opts the forward in too,
// with the sign-safe ReLU pass of gemm_tc_f32.cu behind every ReLU layer.
static GemmCtx make_ctx(const NetWorkspace &W, int precision, int training, int n_points, bool forward) {
  GemmCtx c;
  c.tc = prec_is_tc(precision) && training && n_points >= 128 && W.tc_elems > 0 && W.tc_pa != nullptr &&
         (!forward || precision == PLR2_PREC_BF16X2_ALL);
  c.S.pa = reinterpret_cast<__nv_bfloat16 *>(W.tc_pa);
  c.S.pb = reinterpret_cast<__nv_bfloat16 *>(W.tc_pb);
  c.S.ct = W.tc_ct;
  c.S.elems = W.tc_elems;
  return c;
}
static int gemm_run(const GemmArgs &g, int split_k, const GemmCtx &c, cudaStream_t st) {
  if (c.tc && gemm_tc_eligible(g)) return gemm_tc_f32(g, c.S, st);
  return gemm_f32(g, split_k, st);
}

#define RC(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

static int check_ws(const NetLayout &L, void *ws, size_t bytes, int frames, int n, int training, NetWorkspace &W) {
  if (!ws) { set_error("network: workspace is NULL"); return PLR2_E_NULLPTR; }
  if (reinterpret_cast<uintptr_t>(ws) & 127) { set_error("network: workspace must be 128-byte aligned"); return PLR2_E_ALIGN; }
  W = carve_workspace(L, ws, frames, n, training);
  if (W.floats * sizeof(float) > bytes) {
    set_error("network: workspace too small (%zu < %zu bytes)", bytes, W.floats * sizeof(float));
    return PLR2_E_WORKSPACE;
  }
  return 0;
}

// Input-range flags of a forward live in the first word of its workspace (net_layout.cuh): cleared here, set by
// first_layer_kernel where torch.gather / index_select would have raised, read back by plr2_net_input_status.
static int reset_status(unsigned int *status, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(status, 0, 128, st);
  if (e != cudaSuccess) { set_error("network: status reset: %s", cudaGetErrorString(e)); return (int)e; }
  return 0;
}

// Shared PointNet feature trunk: first layer, conv2/e_conv2, conv5, conv6, per-frame mean.
static int feat_forward(const float *pk, const NetLayout &L, const NetWorkspace &W, const float *out_img,
                        const int64_t *choose, const float *emb_in, const float *x, int F, int N, int hw,
                        float *emb_out, int training, const int64_t *obj, const GemmCtx &cx, cudaStream_t st) {
  const int R = F * N;
  RC(reset_status(W.status, st));
  RC(first_layer(pk, L, out_img, choose, emb_in, x, F, N, hw, W.pf, emb_out, training ? W.embT : nullptr, obj, W.status, st));
  GemmArgs g = gemm_nt(W.pf, 384, pk + L.c2w, 64, W.pf + 128, 384, R, 128, 64);   // conv2
  g.bias = pk + L.c2b; g.relu = 1;
  RC(gemm_run(g, 1, cx, st));
  g = gemm_nt(W.pf + 64, 384, pk + L.e2w, 64, W.pf + 256, 384, R, 128, 64);       // e_conv2
  g.bias = pk + L.e2b; g.relu = 1;
  RC(gemm_run(g, 1, cx, st));
  const float *a5 = L.c5k == 256 ? W.pf + 128 : W.pf;                              // pointfeat_2 | pointfeat_3
  g = gemm_nt(a5, 384, pk + L.c5w, L.c5k, W.x5, 512, R, 512, L.c5k);              // conv5
  g.bias = pk + L.c5b; g.relu = 1;
  RC(gemm_run(g, 1, cx, st));
  g = gemm_nt(W.x5, 512, pk + L.c6w, 512, W.x6, 1024, R, 1024, 512);              // conv6
  g.bias = pk + L.c6b; g.relu = 1;
  RC(gemm_run(g, 1, cx, st));
  RC(segment_colsum(W.x6, 1024, F, N, 1024, 1.0f / (float)N, W.ap, 1024, st));    // AvgPool1d(num_points)
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Tensor-core inference path (PLR2_PREC_BF16X2, no activations kept for backward).
static bool use_tc(int precision, int training, int n_points) {
  return prec_is_tc(precision) && !training && n_points >= 128;   // per-frame epilogues: <= 2 frames per 128-row tile
}

static int check_tc_ws(const NetLayout &L, void *ws, size_t bytes, int frames, int n, TcWorkspace &W) {
  if (!ws) { set_error("network: workspace is NULL"); return PLR2_E_NULLPTR; }
  if (reinterpret_cast<uintptr_t>(ws) & 127) { set_error("network: workspace must be 128-byte aligned"); return PLR2_E_ALIGN; }
  W = carve_tc_workspace(L, ws, frames, n);
  if (W.bytes > bytes) { set_error("network: workspace too small (%zu < %zu bytes)", bytes, W.bytes); return PLR2_E_WORKSPACE; }
  return 0;
}

// first layer, conv2/e_conv2, conv5, conv6 + fused per-frame average pooling -> W.ap [F,1024]
static int feat_forward_tc(const float *pk, const NetLayout &L, const TcWorkspace &W, const float *out_img,
                           const int64_t *choose, const float *emb_in, const float *x, int F, int N, int hw,
                           float *emb_out, bool reuse_emb, const int64_t *obj, cudaStream_t st) {
  const int R = F * N;
  const __nv_bfloat16 *wb = reinterpret_cast<const __nv_bfloat16 *>(pk + L.total);   // split-bf16 weights
  const long long wplane = (long long)L.total;
  __nv_bfloat16 *pf = (__nv_bfloat16 *)W.pf, *x5 = (__nv_bfloat16 *)W.x5;
  // the refiner's trunk is consumed through the pool only (hi-only readers below): no lo plane needed
  RC(reset_status(W.status, st));
  RC(first_layer_bf16(pk, L, out_img, choose, emb_in, x, F, N, hw, pf, L.net == PLR2_NET_REFINER ? 0 : W.pf_plane, emb_out, reuse_emb ? 1 : 0,
                      obj, W.status, st));
  TcGemm g;
  g.A = pf; g.lda = 384; g.a_plane = W.pf_plane; g.W = wb + L.c2w; g.ldw = 64; g.w_plane = wplane;
  g.M = R; g.N = 128; g.K = 64; g.bias = pk + L.c2b; g.relu = 1;
  g.out_mode = TC_OUT_BF16X2; g.Cb = pf + 128; g.ldcb = 384; g.c_plane = W.pf_plane;
  // conv2 and e_conv2 in one launch (batch = branch): x1 -> pf[:,128:256), e1 -> pf[:,256:384)
  g.batch = 2; g.a_col_z = 64; g.b_row_z = (int)((L.e2w - L.c2w) / 64); g.c_col_z = 128;
  g.bias_z = (long long)(L.e2b - L.c2b);
  if (reuse_emb) g.batch = 1;        // pf[:,64:128) and pf[:,256:384) still hold e_conv1 / e_conv2 of the same emb
  // Layers whose output is consumed ONLY through the per-frame average pool (conv5, conv6; in the refiner
  // also conv2/e_conv2) read just the hi plane of their activations (2 MMAs per MAC instead of 3): the
  // bf16 rounding of an activation is unbiased and independent from point to point, so it averages out
  // over the num_points rows of the pool (error ~2^-9/sqrt(N)), while weights keep hi+lo.
  const bool pooled_only = L.net == PLR2_NET_REFINER;
  g.a_hi_only = pooled_only ? 1 : 0;
  if (pooled_only) g.c_plane = 0;
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));
  g = TcGemm();
  g.A = L.c5k == 256 ? pf + 128 : pf; g.lda = 384; g.a_plane = W.pf_plane;
  g.W = wb + L.c5w; g.ldw = L.c5k; g.w_plane = wplane;
  g.M = R; g.N = 512; g.K = L.c5k; g.bias = pk + L.c5b; g.relu = 1;
  g.out_mode = TC_OUT_BF16X2; g.Cb = x5; g.ldcb = 512; g.c_plane = 0;                // hi plane only (see above)
  g.a_hi_only = 1;
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));                                                            // conv5
  g = TcGemm();
  g.A = x5; g.lda = 512; g.a_plane = W.x5_plane; g.W = wb + L.c6w; g.ldw = 512; g.w_plane = wplane;
  g.M = R; g.N = 1024; g.K = 512; g.bias = pk + L.c6b; g.relu = 1;
  g.out_mode = TC_OUT_POOL; g.pool = W.pool; g.rows_per_frame = N;
  g.a_hi_only = 1;
  g.w_hi_only = infer_conv6_nprod() == 1 ? 1 : 0;   // single product: experiment only, see common.cuh
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));                                                            // conv6 + AvgPool partials
  RC(pool_finish(W.pool, F, N, R, 1024, W.ap, (__nv_bfloat16 *)W.apb, W.apb_plane, infer_fp16() ? 1 : 0, st));
  return 0;
}

static int posenet_forward_tc(const float *pk, const NetLayout &L, const TcWorkspace &W, const float *out_img,
                              const int64_t *choose, const float *x, const int64_t *obj, int F, int N, int hw,
                              float *pred_r, float *pred_t, float *pred_c, float *emb, cudaStream_t st) {
  const int R = F * N;
  const __nv_bfloat16 *wb = reinterpret_cast<const __nv_bfloat16 *>(pk + L.total);
  const long long wplane = (long long)L.total;
  __nv_bfloat16 *pf = (__nv_bfloat16 *)W.pf, *h1 = (__nv_bfloat16 *)W.h1, *h2 = (__nv_bfloat16 *)W.h2;
  RC(feat_forward_tc(pk, L, W, out_img, choose, nullptr, x, F, N, hw, emb, false, obj, st));
  // per-frame bias of conv1_{r,t,c}: hb[F,1920] = ap W1[:,384:]^T + b1 (an M = frames GEMM, FP32 out)
  TcGemm g;
  g.A = (const __nv_bfloat16 *)W.apb; g.lda = 1024; g.a_plane = W.apb_plane; g.W = wb + L.h1ap; g.ldw = 1024; g.w_plane = wplane;
  g.M = F; g.N = 1920; g.K = 1024; g.bias = pk + L.h1b; g.out_mode = TC_OUT_F32; g.Cf = W.hb; g.ldc = 1920;
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));
  g = TcGemm();
  g.A = pf; g.lda = 384; g.a_plane = W.pf_plane; g.W = wb + L.h1pf; g.ldw = 384; g.w_plane = wplane;
  g.M = R; g.N = 1920; g.K = 384; g.frame_bias = W.hb; g.rows_per_frame = N; g.ldfb = 1920; g.relu = 1;
  g.out_mode = TC_OUT_BF16X2; g.Cb = h1; g.ldcb = 1920; g.c_plane = W.h1_plane;
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));                                                            // conv1_{r,t,c}
  g = TcGemm();
  g.A = h1; g.lda = 1920; g.a_plane = W.h1_plane; g.W = wb + L.h2w; g.ldw = 640; g.w_plane = wplane;
  g.M = R; g.N = 256; g.K = 640; g.batch = 3; g.a_col_z = 640; g.b_row_z = 256; g.c_col_z = 256; g.bias_z = 256;
  g.bias = pk + L.h2b; g.relu = 1;
  g.out_mode = TC_OUT_BF16X2; g.Cb = h2; g.ldcb = 768; g.c_plane = W.h2_plane;
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));                                                            // conv2_*
  g = TcGemm();
  g.A = h2; g.lda = 768; g.a_plane = W.h2_plane; g.W = wb + L.h3w; g.ldw = 256; g.w_plane = wplane;
  g.M = R; g.N = 128; g.K = 256; g.batch = 3; g.a_col_z = 256; g.b_row_z = 128; g.c_col_z = 128; g.bias_z = 128;
  g.bias = pk + L.h3b; g.relu = 1;
  // conv3_* with conv4_{r,t,c} for the frame's class + sigmoid fused into the epilogue (FP32): the
  // 128-wide h3 activations are never written
  g.out_mode = TC_OUT_HEAD; g.rows_per_frame = N;
  g.w4 = pk + L.h4w; g.b4 = pk + L.h4b; g.obj = obj; g.num_obj = L.O;
  g.pred_r = pred_r; g.pred_t = pred_t; g.pred_c = pred_c;
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));
  return 0;
}

static int refiner_forward_tc(const float *pk, const NetLayout &L, const TcWorkspace &W, const float *x,
                              const float *emb, const int64_t *obj, int F, int N, float *pred_r, float *pred_t,
                              bool reuse_emb, cudaStream_t st) {
  RC(feat_forward_tc(pk, L, W, nullptr, nullptr, emb, x, F, N, 0, nullptr, reuse_emb, obj, st));
  // Linear 1024->512 (x2 heads) as one M = frames tensor-core GEMM; 512->128 (x2) and 128->7 for the frame's
  // class in FP32 SIMT
  const __nv_bfloat16 *wb = reinterpret_cast<const __nv_bfloat16 *>(pk + L.total);
  TcGemm g;
  g.A = (const __nv_bfloat16 *)W.apb; g.lda = 1024; g.a_plane = W.apb_plane; g.W = wb + L.h1ap; g.ldw = 1024;
  g.w_plane = (long long)L.total;
  g.M = F; g.N = 1024; g.K = 1024; g.bias = pk + L.h1b; g.relu = 1; g.out_mode = TC_OUT_F32; g.Cf = W.hb; g.ldc = 1024;
  g.fp16 = infer_fp16(); RC(gemm_bf16x2(g, st));
  RC(frame_linear(W.hb, 1024, pk + L.h2w, 512, pk + L.h2b, W.h3, 256, F, 128, 512, 1, 2, 512, 128 * 512, 128, 128, st));
  RC(head_out(W.h3, 256, pk + L.h4w, pk + L.h4b, obj, L.O, F, 1, pred_r, pred_t, nullptr, st));
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Backward helpers.  Every dX is written IN PLACE over the activation whose ReLU it passes through
// (the mask is read and the gradient written by the same thread), after the weight gradient that
// still needs that activation has been issued.
static int auto_split(int M, int N, int batch, int K) {
  const long long tiles = (long long)((M + 127) / 128) * ((N + 127) / 128) * batch;
  long long s = (2 * 148 + tiles - 1) / tiles;
  const long long smax = K / 256 > 0 ? K / 256 : 1;
  if (s > smax) s = smax;
  return (int)(s < 1 ? 1 : s);
}
static int wgrad(GemmArgs g, const GemmCtx &cx, cudaStream_t st) { return gemm_run(g, auto_split(g.M, g.N, g.batch, g.K), cx, st); }
// db[c] = sum over rows of dY[r][c].  Tall matrices are summed in two fixed-order stages (row slices into `red`,
// then the slices), so that the first stage fills the GPU instead of running ncols/32 blocks.
static int bgrad(const float *dY, long long ld, int rows, int ncols, float *db, float *red, cudaStream_t st) {
  const int per = 256;                                    // rows per slice (a multiple of the kernel's 8 row lanes)
  int slices = (rows + per - 1) / per;
  if (red == nullptr || slices < 4 || ncols > BGRAD_MAX_COLS) return segment_colsum(dY, ld, 1, rows, ncols, 1.0f, db, ncols, st);
  int rows_per = per;
  if (slices > BGRAD_SLICES) { slices = BGRAD_SLICES; rows_per = ((rows + slices - 1) / slices + 7) / 8 * 8; slices = (rows + rows_per - 1) / rows_per; }
  RC(segment_colsum_ragged(dY, ld, slices, rows_per, rows, ncols, red, ncols, st));
  return segment_colsum(red, ncols, 1, slices, ncols, 1.0f, db, ncols, st);
}

// Weight and bias gradient of one layer (or of the stacked r|t|c heads): dW += dY^T X, db = column sums of dY.
// On the tensor-core path the pass that splits dY leaves per-tile column sums in W.red, so only their sum is left
// to do; otherwise dY is summed in two stages as before.
static int wbgrad(GemmArgs g, const float *dY, long long ld, int rows, int ncols, float *db, const NetWorkspace &W,
                  const GemmCtx &cx, cudaStream_t st) {
  int tiles = 0;
  g.colsum_part = W.red; g.colsum_cap = (long long)W.red_cap; g.colsum_tiles = &tiles;
  RC(wgrad(g, cx, st));
  if (tiles > 0) return colsum_tall(W.red, ncols, tiles, ncols, db, st);
  return bgrad(dY, ld, rows, ncols, db, W.red, st);
}

// From W.dap (d loss / d pooled feature) and W.dpf (head contributions to d pf, or undefined when
// !dpf_has_heads) down to the first layer.  Leaves d e1 in W.dpf[:,64:128] and d emb in W.embT.
static int feat_backward(const float *pk, float *gp, const NetLayout &L, const NetWorkspace &W, const float *x,
                         int F, int N, bool dpf_has_heads, const GemmCtx &cx, cudaStream_t st) {
  const int R = F * N;
  // AvgPool1d backward through conv6's ReLU: dx6 = (x6 > 0) * dap[frame] / N     (in place)
  RC(segment_bcast_mask(W.dap, W.x6, (size_t)R, N, 1024, 1.0f / (float)N, W.x6, st));
  RC(wbgrad(gemm_tn(W.x6, 1024, W.x5, 512, gp + L.c6w, 512, 1024, 512, R), W.x6, 1024, R, 1024, gp + L.c6b, W, cx, st));
  GemmArgs g = gemm_nn(W.x6, 1024, pk + L.c6w, 512, W.x5, 512, R, 512, 1024);     // dx5 (in place)
  g.relu_mask = W.x5; g.ldmask = 512; g.reuse_a_planes = 1;
  RC(gemm_run(g, 1, cx, st));
  const float *a5 = L.c5k == 256 ? W.pf + 128 : W.pf;
  RC(wbgrad(gemm_tn(W.x5, 512, a5, 384, gp + L.c5w, L.c5k, 512, L.c5k, R), W.x5, 512, R, 512, gp + L.c5b, W, cx, st));
  const int off2 = L.c5k == 256 ? 0 : 128;    // column of pointfeat_2 inside conv5's input
  if (L.c5k == 384) {                         // refiner: conv5 also reads pointfeat_1
    g = gemm_nn(W.x5, 512, pk + L.c5w, 384, W.dpf, 384, R, 128, 512);
    g.accumulate = dpf_has_heads ? 1 : 0; g.reuse_a_planes = 1;
    RC(gemm_run(g, 1, cx, st));
  }
  g = gemm_nn(W.x5, 512, pk + L.c5w + off2, L.c5k, W.dpf + 128, 384, R, 256, 512);  // d pointfeat_2
  g.accumulate = dpf_has_heads ? 1 : 0; g.reuse_a_planes = 1;
  g.relu_mask = W.pf + 128; g.ldmask = 384;
  RC(gemm_run(g, 1, cx, st));
  // conv2 / e_conv2
  for (int e = 0; e < 2; ++e) {
    const int cin = e * 64, cout = 128 + e * 128;
    const size_t wo = e ? L.e2w : L.c2w, bo = e ? L.e2b : L.c2b;
    RC(wbgrad(gemm_tn(W.dpf + cout, 384, W.pf + cin, 384, gp + wo, 64, 128, 64, R), W.dpf + cout, 384, R, 128, gp + bo, W, cx, st));
    g = gemm_nn(W.dpf + cout, 384, pk + wo, 64, W.dpf + cin, 384, R, 64, 128);
    g.accumulate = 1; g.relu_mask = W.pf + cin; g.ldmask = 384; g.reuse_a_planes = 1;
    RC(gemm_run(g, 1, cx, st));
  }
  // conv1 (3 -> 64) and e_conv1 (32 -> 64)
  RC(wgrad(gemm_tn(W.dpf, 384, x, 3, gp + L.c1w, 4, 64, 3, R), cx, st));
  RC(bgrad(W.dpf, 384, R, 64, gp + L.c1b, W.red, st));
  RC(wbgrad(gemm_tn(W.dpf + 64, 384, W.embT, 32, gp + L.e1w, 32, 64, 32, R), W.dpf + 64, 384, R, 64, gp + L.e1b, W, cx, st));
  return 0;
}

}  // namespace plr2

using namespace plr2;

static thread_local cudaEvent_t g_stage_event = nullptr;

extern "C" int plr2_set_backward_stage_event(void *event) {
  g_stage_event = (cudaEvent_t)event;
  return 0;
}

extern "C" int plr2_net_grad_regions(int net, int num_obj, size_t *early_begin, size_t *early_end) {
  if (!early_begin || !early_end || num_obj <= 0 || (net != PLR2_NET_POSENET && net != PLR2_NET_REFINER)) {
    set_error("plr2_net_grad_regions: bad arguments");
    return PLR2_E_SHAPE;
  }
  const NetLayout L = make_layout(net, num_obj);
  *early_begin = L.h1pf;      // head weights: conv1_*..conv4_* (PoseNet), conv1_*..conv3_* (refiner)
  *early_end = L.c1b;         // biases of all layers follow the weights
  return 0;
}

extern "C" int plr2_net_input_status(const void *workspace, unsigned int *flags, void *stream) {
  if (!workspace || !flags) { set_error("plr2_net_input_status: null pointer"); return PLR2_E_NULLPTR; }
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyAsync(flags, workspace, sizeof(unsigned int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { set_error("plr2_net_input_status: %s", cudaGetErrorString(e)); return (int)e; }
  return 0;
}

extern "C" size_t plr2_net_packed_floats(int net, int num_obj) { return 2 * make_layout(net, num_obj).total; }

extern "C" size_t plr2_net_workspace_bytes(int net, int frames, int n_points, int num_obj, int training) {
  const NetLayout L = make_layout(net, num_obj);
  const size_t a = carve_workspace(L, nullptr, frames, n_points, training).floats * sizeof(float);
  const size_t b = carve_tc_workspace(L, nullptr, frames, n_points).bytes;
  return a > b ? a : b;
}

extern "C" int plr2_net_pack(int net, const plr2_param_table *params, int num_obj, float *packed, int split_bf16,
                             void *stream) {
  if (!params || !packed) { set_error("plr2_net_pack: null pointer"); return PLR2_E_NULLPTR; }
  if (num_obj <= 0 || (net != PLR2_NET_POSENET && net != PLR2_NET_REFINER)) { set_error("plr2_net_pack: bad net/num_obj"); return PLR2_E_SHAPE; }
  return net_copy_params(make_layout(net, num_obj), params, packed, true, split_bf16, (cudaStream_t)stream);
}

extern "C" int plr2_net_unpack_grads(int net, const float *packed_grad, const plr2_param_table *grads, int num_obj,
                                     void *stream) {
  if (!grads || !packed_grad) { set_error("plr2_net_unpack_grads: null pointer"); return PLR2_E_NULLPTR; }
  if (num_obj <= 0 || (net != PLR2_NET_POSENET && net != PLR2_NET_REFINER)) { set_error("plr2_net_unpack_grads: bad net/num_obj"); return PLR2_E_SHAPE; }
  return net_copy_params(make_layout(net, num_obj), grads, const_cast<float *>(packed_grad), false, 0,
                         (cudaStream_t)stream);
}

extern "C" int plr2_net_accumulate_grads(int net, const float *packed_grad, const plr2_param_table *grads, int num_obj,
                                         void *stream) {
  if (!grads || !packed_grad) { set_error("plr2_net_accumulate_grads: null pointer"); return PLR2_E_NULLPTR; }
  if (num_obj <= 0 || (net != PLR2_NET_POSENET && net != PLR2_NET_REFINER)) { set_error("plr2_net_accumulate_grads: bad net/num_obj"); return PLR2_E_SHAPE; }
  return net_copy_params(make_layout(net, num_obj), grads, const_cast<float *>(packed_grad), false, 0,
                         (cudaStream_t)stream, 1);
}

extern "C" int plr2_posenet_forward(const float *packed, const float *out_img, const int64_t *choose,
                                    const float *x, const int64_t *obj, int frames, int n_points, int hw,
                                    int num_obj, void *workspace, size_t workspace_bytes, float *pred_r,
                                    float *pred_t, float *pred_c, float *emb, int training, int precision,
                                    void *stream) {
  if (frames < 0 || n_points <= 0 || hw <= 0 || num_obj <= 0) { set_error("plr2_posenet_forward: bad shape"); return PLR2_E_SHAPE; }
  if (frames == 0) return 0;
  if (!packed || !out_img || !choose || !x || !obj || !pred_r || !pred_t || !pred_c) { set_error("plr2_posenet_forward: null pointer"); return PLR2_E_NULLPTR; }
  if ((long long)frames * n_points > 0x7fffffffLL / 2048) { set_error("plr2_posenet_forward: too many rows per call"); return PLR2_E_SHAPE; }
  const NetLayout L = make_layout(PLR2_NET_POSENET, num_obj);
  cudaStream_t st = (cudaStream_t)stream;
  if (use_tc(precision, training, n_points)) {
    TcWorkspace T;
    RC(check_tc_ws(L, workspace, workspace_bytes, frames, n_points, T));
    return posenet_forward_tc(packed, L, T, out_img, choose, x, obj, frames, n_points, hw, pred_r, pred_t, pred_c, emb, st);
  }
  NetWorkspace W;
  RC(check_ws(L, workspace, workspace_bytes, frames, n_points, training, W));
  const GemmCtx cx = make_ctx(W, precision, training, n_points, true);
  const int F = frames, N = n_points, R = F * N;
  const float *pk = packed;
  RC(feat_forward(pk, L, W, out_img, choose, nullptr, x, F, N, hw, emb, training, obj, cx, st));
  // per-frame bias of the three conv1 heads: hb[F,1920] = ap W1[:,384:]^T + b1   (FP32 always)
  GemmArgs g = gemm_nt(W.ap, 1024, pk + L.h1ap, 1024, W.hb, 1920, F, 1920, 1024);
  g.bias = pk + L.h1b;
  RC(gemm_run(g, 1, cx, st));
  // conv1_{r,t,c} on the per-point 384 channels, all three heads in one launch
  g = gemm_nt(W.pf, 384, pk + L.h1pf, 384, W.h1, 1920, R, 1920, 384);
  g.frame_bias = W.hb; g.rows_per_frame = N; g.ldfb = 1920; g.relu = 1;
  RC(gemm_run(g, 1, cx, st));
  // conv2_* : block diagonal, batch = head
  g = gemm_nt(W.h1, 1920, pk + L.h2w, 640, W.h2, 768, R, 256, 640);
  g.bias = pk + L.h2b; g.relu = 1; g.batch = 3; g.zA = 640; g.zB = 256 * 640; g.zC = 256; g.zbias = 256;
  RC(gemm_run(g, 1, cx, st));
  // conv3_*
  g = gemm_nt(W.h2, 768, pk + L.h3w, 256, W.h3, 384, R, 128, 256);
  g.bias = pk + L.h3b; g.relu = 1; g.batch = 3; g.zA = 256; g.zB = 128 * 256; g.zC = 128; g.zbias = 128;
  RC(gemm_run(g, 1, cx, st));
  // conv4_* for every class (the class of each frame is selected afterwards)
  size_t row = 0;
  for (int h = 0; h < 3; ++h) {
    g = gemm_nt(W.h3 + h * 128, 384, pk + L.h4w + row * 128, 128, W.out + L.oc[h], L.ldo, R, L.on[h], 128);
    g.bias = pk + L.h4b + row;
    RC(gemm_run(g, 1, cx, st));
    row += L.on[h];
  }
  RC(select_outputs(W.out, L, obj, N, (size_t)R, pred_r, pred_t, pred_c, st));
  return 0;
}

extern "C" int plr2_refiner_forward(const float *packed, const float *x, const float *emb, const int64_t *obj,
                                    int frames, int n_points, int num_obj, void *workspace,
                                    size_t workspace_bytes, float *pred_r, float *pred_t, int training,
                                    int precision, void *stream) {
  if (frames < 0 || n_points <= 0 || num_obj <= 0) { set_error("plr2_refiner_forward: bad shape"); return PLR2_E_SHAPE; }
  if (frames == 0) return 0;
  if (!packed || !x || !emb || !obj || !pred_r || !pred_t) { set_error("plr2_refiner_forward: null pointer"); return PLR2_E_NULLPTR; }
  if ((long long)frames * n_points > 0x7fffffffLL / 2048) { set_error("plr2_refiner_forward: too many rows per call"); return PLR2_E_SHAPE; }
  const NetLayout L = make_layout(PLR2_NET_REFINER, num_obj);
  cudaStream_t st = (cudaStream_t)stream;
  const bool next_iter = training == PLR2_REFINER_NEXT_ITER;    // inference, emb branch already in the workspace
  if (next_iter) training = 0;
  if (use_tc(precision, training, n_points)) {
    TcWorkspace T;
    RC(check_tc_ws(L, workspace, workspace_bytes, frames, n_points, T));
    return refiner_forward_tc(packed, L, T, x, emb, obj, frames, n_points, pred_r, pred_t, next_iter, st);
  }
  NetWorkspace W;
  RC(check_ws(L, workspace, workspace_bytes, frames, n_points, training, W));
  const GemmCtx cx = make_ctx(W, precision, training, n_points, true);
  const int F = frames, N = n_points;
  const float *pk = packed;
  RC(feat_forward(pk, L, W, nullptr, nullptr, emb, x, F, N, 0, nullptr, training, obj, cx, st));
  // Linear 1024 -> 512 (x2 heads), 512 -> 128, 128 -> num_obj*{4,3}: M = frames, FP32
  GemmArgs g = gemm_nt(W.ap, 1024, pk + L.h1ap, 1024, W.h1, 1024, F, 1024, 1024);
  g.bias = pk + L.h1b; g.relu = 1;
  RC(gemm_run(g, 1, cx, st));
  g = gemm_nt(W.h1, 1024, pk + L.h2w, 512, W.h2, 256, F, 128, 512);
  g.bias = pk + L.h2b; g.relu = 1; g.batch = 2; g.zA = 512; g.zB = 128 * 512; g.zC = 128; g.zbias = 128;
  RC(gemm_run(g, 1, cx, st));
  size_t row = 0;
  for (int h = 0; h < 2; ++h) {
    g = gemm_nt(W.h2 + h * 128, 256, pk + L.h4w + row * 128, 128, W.out + L.oc[h], L.ldo, F, L.on[h], 128);
    g.bias = pk + L.h4b + row;
    RC(gemm_run(g, 1, cx, st));
    row += L.on[h];
  }
  RC(select_outputs(W.out, L, obj, 1, (size_t)F, pred_r, pred_t, nullptr, st));
  return 0;
}

extern "C" int plr2_posenet_backward(const float *packed, const int64_t *choose, const float *x,
                                     const int64_t *obj, int frames, int n_points, int hw, int num_obj,
                                     void *workspace, size_t workspace_bytes, const float *g_pred_r,
                                     const float *g_pred_t, const float *g_pred_c, float *packed_grad,
                                     float *g_out_img, int precision, void *stream) {
  if (frames <= 0 || n_points <= 0 || num_obj <= 0) { set_error("plr2_posenet_backward: bad shape"); return PLR2_E_SHAPE; }
  if (!packed || !x || !obj || !g_pred_r || !g_pred_t || !g_pred_c || !packed_grad || (g_out_img && !choose)) {
    set_error("plr2_posenet_backward: null pointer"); return PLR2_E_NULLPTR;
  }
  const NetLayout L = make_layout(PLR2_NET_POSENET, num_obj);
  NetWorkspace W;
  RC(check_ws(L, workspace, workspace_bytes, frames, n_points, 1, W));
  const GemmCtx cx = make_ctx(W, precision, 1, n_points, false);
  cudaStream_t st = (cudaStream_t)stream;
  const int F = frames, N = n_points, R = F * N;
  const float *pk = packed;
  float *gp = packed_grad;
  cudaError_t e = cudaMemsetAsync(gp, 0, L.total * sizeof(float), st);
  if (e != cudaSuccess) { set_error("plr2_posenet_backward: memset: %s", cudaGetErrorString(e)); return (int)e; }
  RC(select_outputs_bwd(W.out, L, obj, N, (size_t)R, g_pred_r, g_pred_t, g_pred_c, st));
  size_t row = 0;
  for (int h = 0; h < 3; ++h) {               // conv4_*
    RC(wbgrad(gemm_tn(W.out + L.oc[h], L.ldo, W.h3 + h * 128, 384, gp + L.h4w + row * 128, 128, L.on[h], 128, R),
              W.out + L.oc[h], L.ldo, R, L.on[h], gp + L.h4b + row, W, cx, st));
    GemmArgs g = gemm_nn(W.out + L.oc[h], L.ldo, pk + L.h4w + row * 128, 128, W.h3 + h * 128, 384, R, 128, L.on[h]);
    g.relu_mask = W.h3 + h * 128; g.ldmask = 384; g.reuse_a_planes = 1;
    RC(gemm_run(g, 1, cx, st));
    row += L.on[h];
  }
  // conv3_*
  GemmArgs g = gemm_tn(W.h3, 384, W.h2, 768, gp + L.h3w, 256, 128, 256, R);
  g.batch = 3; g.zA = 128; g.zB = 256; g.zC = 128 * 256;
  RC(wbgrad(g, W.h3, 384, R, 384, gp + L.h3b, W, cx, st));
  g = gemm_nn(W.h3, 384, pk + L.h3w, 256, W.h2, 768, R, 256, 128);
  g.batch = 3; g.zA = 128; g.zB = 128 * 256; g.zC = 256; g.relu_mask = W.h2; g.ldmask = 768; g.zmask = 256;
  g.reuse_a_planes = 1;
  RC(gemm_run(g, 1, cx, st));
  // conv2_*
  g = gemm_tn(W.h2, 768, W.h1, 1920, gp + L.h2w, 640, 256, 640, R);
  g.batch = 3; g.zA = 256; g.zB = 640; g.zC = 256 * 640;
  RC(wbgrad(g, W.h2, 768, R, 768, gp + L.h2b, W, cx, st));
  g = gemm_nn(W.h2, 768, pk + L.h2w, 640, W.h1, 1920, R, 640, 256);
  g.batch = 3; g.zA = 256; g.zB = 256 * 640; g.zC = 640; g.relu_mask = W.h1; g.ldmask = 1920; g.zmask = 640;
  g.reuse_a_planes = 1;
  RC(gemm_run(g, 1, cx, st));
  // conv1_*: per-point part, per-frame (pooled) part, bias
  RC(wgrad(gemm_tn(W.h1, 1920, W.pf, 384, gp + L.h1pf, 384, 1920, 384, R), cx, st));
  {  // d pf (head part) straight after the weight gradient: both read the same split of d h1
    GemmArgs gd = gemm_nn(W.h1, 1920, pk + L.h1pf, 384, W.dpf, 384, R, 384, 1920);
    gd.reuse_a_planes = 1;
    RC(gemm_run(gd, 1, cx, st));
  }
  RC(segment_colsum(W.h1, 1920, F, N, 1920, 1.0f, W.dhb, 1920, st));
  RC(bgrad(W.dhb, 1920, F, 1920, gp + L.h1b, W.red, st));
  RC(gemm_run(gemm_tn(W.dhb, 1920, W.ap, 1024, gp + L.h1ap, 1024, 1920, 1024, F), 1, cx, st));
  {  // d ap = d hb . W1[:,384:]: M = frames rows only -> split K over CTAs (atomic accumulation onto zeros)
    cudaError_t e = cudaMemsetAsync(W.dap, 0, (size_t)F * 1024 * sizeof(float), st);
    if (e != cudaSuccess) { set_error("posenet_backward: memset: %s", cudaGetErrorString(e)); return (int)e; }
    GemmArgs ga = gemm_nn(W.dhb, 1920, pk + L.h1ap, 1024, W.dap, 1024, F, 1024, 1920);
    ga.accumulate = 1;
    RC(gemm_run(ga, auto_split(F, 1024, 1, 1920), cx, st));
  }
  // every head-weight gradient (packed region [h1pf, c1b)) is final here: a caller that overlaps its gradient
  // all-reduce with the trunk backward waits on this event (plr2_set_backward_stage_event, one-shot)
  if (g_stage_event != nullptr) {
    cudaError_t e = cudaEventRecord(g_stage_event, st);
    g_stage_event = nullptr;
    if (e != cudaSuccess) { set_error("plr2_posenet_backward: stage event: %s", cudaGetErrorString(e)); return (int)e; }
  }
  RC(feat_backward(pk, gp, L, W, x, F, N, true, cx, st));
  if (g_out_img != nullptr) {                 // d emb = d e1 . We1, scattered back to the pixels
    RC(gemm_run(gemm_nn(W.dpf + 64, 384, pk + L.e1w, 32, W.embT, 32, R, 32, 64), 1, cx, st));
    RC(scatter_emb_grad(W.embT, choose, F, N, hw, g_out_img, st));
  }
  return 0;
}

extern "C" int plr2_refiner_backward(const float *packed, const float *x, const float *emb, const int64_t *obj,
                                     int frames, int n_points, int num_obj, void *workspace,
                                     size_t workspace_bytes, const float *g_pred_r, const float *g_pred_t,
                                     float *packed_grad, int precision, void *stream) {
  (void)emb;
  if (frames <= 0 || n_points <= 0 || num_obj <= 0) { set_error("plr2_refiner_backward: bad shape"); return PLR2_E_SHAPE; }
  if (!packed || !x || !obj || !g_pred_r || !g_pred_t || !packed_grad) { set_error("plr2_refiner_backward: null pointer"); return PLR2_E_NULLPTR; }
  const NetLayout L = make_layout(PLR2_NET_REFINER, num_obj);
  NetWorkspace W;
  RC(check_ws(L, workspace, workspace_bytes, frames, n_points, 1, W));
  const GemmCtx cx = make_ctx(W, precision, 1, n_points, false);
  cudaStream_t st = (cudaStream_t)stream;
  const int F = frames, N = n_points;
  const float *pk = packed;
  float *gp = packed_grad;
  cudaError_t e = cudaMemsetAsync(gp, 0, L.total * sizeof(float), st);
  if (e != cudaSuccess) { set_error("plr2_refiner_backward: memset: %s", cudaGetErrorString(e)); return (int)e; }
  RC(select_outputs_bwd(W.out, L, obj, 1, (size_t)F, g_pred_r, g_pred_t, nullptr, st));
  size_t row = 0;
  for (int h = 0; h < 2; ++h) {               // conv3_{r,t}
    RC(gemm_run(gemm_tn(W.out + L.oc[h], L.ldo, W.h2 + h * 128, 256, gp + L.h4w + row * 128, 128, L.on[h], 128, F), 1, cx, st));
    RC(bgrad(W.out + L.oc[h], L.ldo, F, L.on[h], gp + L.h4b + row, W.red, st));
    GemmArgs g = gemm_nn(W.out + L.oc[h], L.ldo, pk + L.h4w + row * 128, 128, W.h2 + h * 128, 256, F, 128, L.on[h]);
    g.relu_mask = W.h2 + h * 128; g.ldmask = 256;
    RC(gemm_run(g, 1, cx, st));
    row += L.on[h];
  }
  GemmArgs g = gemm_tn(W.h2, 256, W.h1, 1024, gp + L.h2w, 512, 128, 512, F);      // conv2_{r,t}
  g.batch = 2; g.zA = 128; g.zB = 512; g.zC = 128 * 512;
  RC(gemm_run(g, 1, cx, st));
  RC(bgrad(W.h2, 256, F, 256, gp + L.h2b, W.red, st));
  g = gemm_nn(W.h2, 256, pk + L.h2w, 512, W.h1, 1024, F, 512, 128);
  g.batch = 2; g.zA = 128; g.zB = 128 * 512; g.zC = 512; g.relu_mask = W.h1; g.ldmask = 1024; g.zmask = 512;
  RC(gemm_run(g, 1, cx, st));
  RC(gemm_run(gemm_tn(W.h1, 1024, W.ap, 1024, gp + L.h1ap, 1024, 1024, 1024, F), 1, cx, st));   // conv1_{r,t}
  RC(bgrad(W.h1, 1024, F, 1024, gp + L.h1b, W.red, st));
  RC(gemm_run(gemm_nn(W.h1, 1024, pk + L.h1ap, 1024, W.dap, 1024, F, 1024, 1024), 1, cx, st));
  RC(feat_backward(pk, gp, L, W, x, F, N, false, cx, st));
  return 0;
}
