/*
 * plr2_b200.h — C-ABI of the B200-native DenseFusion dense-fusion + pose-loss hot path.
 *
 * This is the drop-in boundary. Every entry point takes plain device pointers,
 * sizes and a cudaStream_t (passed as void*), returns an int status and never throws:
 *     0  ok      <0  invalid argument (PLR2_E_*)      >0  a cudaError_t value
 * No entry point allocates: scratch is passed by the caller (see the *_workspace_bytes queries).
 * All launches are asynchronous on `stream`; nothing here synchronises the host.
 *
 * Reference interfaces replaced (paths relative to the reference repo root):
 *   plr2_knn_f32            <- knn(ref, query, idx)            lib/knn/src/knn.h:12-61, src/vision.cpp:3-5,
 *                              knn_device                       lib/knn/src/cuda/knn.cu:217-263
 *   plr2_pose_loss_fwd      <- loss_calculation                 lib/loss.py:12-49 (+ ADD-S branch :39-46)
 *   plr2_pose_loss_bwd      <- autograd of the above            (torch autograd in the reference)
 *   plr2_pose_select        <- best-vote select + re-centre     lib/loss.py:52-69
 *   plr2_refine_loss_fwd/bwd<- loss_calculation (refiner)       lib/loss_refiner.py:12-62
 *   plr2_gemm_f32 / plr2_gemm_bf16x2 <- Conv1d(k=1)/Linear library calls  lib/network.py:54-63,107-121,152-163,193-200
 *   plr2_posenet_* / plr2_refiner_*  <- PoseNet.forward / PoseRefineNet.forward  lib/network.py:95-132,187-206
 */
#ifndef PLR2_B200_H
#define PLR2_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PLR2_ABI_VERSION 1

#define PLR2_E_NULLPTR   (-1)
#define PLR2_E_SHAPE     (-2)
#define PLR2_E_UNSUPPORTED (-3)
#define PLR2_E_WORKSPACE (-4)
#define PLR2_E_ALIGN     (-5)

/* KNN arithmetic modes.  EXACT reproduces the reference CUDA kernel's arithmetic bit for bit
 * (sum over d of fma(a-b, a-b, acc), lib/knn/src/cuda/knn.cu:88-91, strict '<' so the lowest
 * reference index wins ties, :158).  FAST ranks by |r|^2 - 2 q.r on centroid-centred coordinates
 * (3 FMAs/pair); it returns the same index wherever best and second-best differ by >1e-6. */
#define PLR2_KNN_EXACT 0
#define PLR2_KNN_FAST  1

int         plr2_abi_version(void);
const char *plr2_last_error_string(void);   /* thread-local, never NULL */
int         plr2_device_sm_count(void);     /* multiProcessorCount of the current device, or <0 */
long long   plr2_launch_count(void);        /* kernels launched by this library so far (this process) */
/* Bracket every tensor-core GEMM launch with CUDA events; end() synchronises the device and returns
 * the summed kernel time (ms), launch count and algorithmic FLOPs (2 M N K) of those launches. */
int         plr2_profile_begin(void);
int         plr2_profile_end(double *ms, long long *launches, double *flops);
double      plr2_profile_executed_flops(void);  /* bf16 MMA FLOPs issued in the last window (2-3x the algorithmic) */

/* ---- (4) brute-force nearest neighbours ---------------------------------------------------
 * ref   [batch][dim][n_ref]   f32, contiguous        (lib/knn/src/knn.h:33)
 * query [batch][dim][n_query] f32, contiguous
 * idx   [batch][k][n_query]   i64, 1-BASED           (knn.cu:128,147,171)
 * dist  [batch][k][n_query]   f32 squared distances of the winners, or NULL
 * dim==3,k==1 runs the tiled FP32x2 kernel; any other (dim,k<=64) runs the generic kernel. */
size_t plr2_knn_workspace_bytes(int batch, int dim, int n_ref, int n_query, int k);
int plr2_knn_f32(const float *ref, const float *query, int64_t *idx, float *dist,
                 int batch, int dim, int n_ref, int n_query, int k, int mode,
                 void *workspace, size_t workspace_bytes, void *stream);
/* Host-side work split of the dim==3,k==1 kernel (replaces the grid arithmetic of knn_device,
 * lib/knn/src/cuda/knn.cu:217-263): every batch element's queries are cut into *nsplit equal parts of
 * *part queries (a multiple of 128), one CTA each, chosen so that batch * nsplit CTAs of at most 9 queries
 * per lane run in as few full waves of `resident_ctas` as possible.  Pure host arithmetic (no GPU). */
int plr2_knn_plan(int batch, int n_query, int resident_ctas, int *nsplit, int *part);

/* ---- (3) fused quaternion->R, mesh transform, ADD / ADD-S, confidence-weighted loss -------
 * One call handles `frames` independent frames (the reference only ever has frames==1).
 *   pred_r [F][N][4] (w,x,y,z, un-normalised)   pred_t [F][N][3]   pred_c [F][N]
 *   target [F][M][3]   model_points [F][M][3]   points [F][N][3]
 *   obj    [F] i64 class ids;  sym_lut [num_obj] u8 (1 = symmetric class, lib/loss.py:40)
 * ADD-S (nearest target point) is used iff sym_lut[obj[f]] && !refine  (lib/loss.py:39-40).
 * Outputs:
 *   dis    [F][N]   mean_j || R_n m_j + p_n + t_n - tgt_nj ||      (lib/loss.py:48)
 *   loss   [F]      mean_n (dis_n c_n - w log c_n)                 (lib/loss.py:49)
 *   jac_r  [F][N][4], jac_t [F][N][3]   d dis_n / d pred_r[n], d dis_n / d pred_t[n]
 *                   (NN index held fixed, lib/loss.py:44 `.detach()`); may be NULL (inference)
 */
int plr2_pose_loss_fwd(const float *pred_r, const float *pred_t, const float *pred_c,
                       const float *target, const float *model_points, const float *points,
                       const int64_t *obj, const uint8_t *sym_lut, int num_obj,
                       int frames, int n_points, int n_mesh, float w, int refine, int knn_mode,
                       float *dis, float *loss, float *jac_r, float *jac_t, void *stream);

/* Backward of the above given upstream grads of `loss` [F] and of the selected `dis_sel` [F]
 * (either may be NULL = zero).  which_max [F] i32 comes from plr2_pose_select.
 *   g_pred_r [F][N][4], g_pred_t [F][N][3], g_pred_c [F][N] */
int plr2_pose_loss_bwd(const float *g_loss, const float *g_dis_sel, const int32_t *which_max,
                       const float *pred_c, const float *dis, const float *jac_r, const float *jac_t,
                       int frames, int n_points, float w,
                       float *g_pred_r, float *g_pred_t, float *g_pred_c, void *stream);

/* ---- (5) best-vote select + re-centre (lib/loss.py:52-69) ---------------------------------
 *   which_max [F] i32 = argmax_n pred_c (lowest n on ties)
 *   dis_sel   [F]     = dis[f][which_max]
 *   new_points[F][N][3] = (points - t*) R*      new_target [F][M][3] = (target - t*) R*
 *   pose_out  [F][12]   = R* row-major (9) then t* (3); may be NULL */
int plr2_pose_select(const float *pred_r, const float *pred_t, const float *pred_c,
                     const float *dis, const float *target, const float *points,
                     int frames, int n_points, int n_mesh,
                     int32_t *which_max, float *dis_sel, float *new_points, float *new_target,
                     float *pose_out, void *stream);

/* ---- refiner loss (lib/loss_refiner.py:12-62): one pose per frame -------------------------
 *   pred_r [F][4], pred_t [F][3]; target/model_points [F][M][3]; points [F][N][3]
 *   ADD-S iff sym_lut[obj[f]] (always, lib/loss_refiner.py:40)
 *   dis [F]; new_points [F][N][3]; new_target [F][M][3]; jac_r [F][4], jac_t [F][3] (nullable)
 *   pose_out [F][12] nullable.  Needs n_mesh <= 8192.
 *   workspace: plr2_refine_loss_workspace_bytes(frames, n_mesh) bytes, 16-byte aligned, contents
 *   irrelevant on entry (per-frame arrival counters + per-CTA partial sums). */
size_t plr2_refine_loss_workspace_bytes(int frames, int n_mesh);
int plr2_refine_loss_fwd(const float *pred_r, const float *pred_t,
                         const float *target, const float *model_points, const float *points,
                         const int64_t *obj, const uint8_t *sym_lut, int num_obj,
                         int frames, int n_points, int n_mesh, int knn_mode,
                         float *dis, float *new_points, float *new_target,
                         float *jac_r, float *jac_t, float *pose_out,
                         void *workspace, size_t workspace_bytes, void *stream);

/* The same refiner loss for callers that cannot provide (or allocate) scratch memory: code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Arguments and results are those of plr2_refine_loss_fwd without the workspace; the arrival counters
 * and partial sums live in a device array inside the library (plr2_refine_loss_static_scratch_bytes() = 427 008 bytes,
 * held by every GPU that loads the library until exit), so the call allocates nothing.
 * Limits: frames <= PLR2_REFINE_LOSS_STATIC_MAX_FRAMES and n_mesh <= 8192; anything larger returns
 * PLR2_E_UNSUPPORTED with a message naming plr2_refine_loss_fwd.  ALL calls on a device share the one array: calls that
 * overlap on the GPU (different streams, concurrent graph branches, several host threads) race on it - serialise
 * them, or use plr2_refine_loss_fwd with one workspace per concurrent call.  Calls on one stream are safe back to back
 * (each call re-clears the counters on its stream before its launch). */
#define PLR2_REFINE_LOSS_STATIC_MAX_FRAMES 256
size_t plr2_refine_loss_static_scratch_bytes(void);
int plr2_refine_loss_fwd_static(const float *pred_r, const float *pred_t,
                                const float *target, const float *model_points, const float *points,
                                const int64_t *obj, const uint8_t *sym_lut, int num_obj,
                                int frames, int n_points, int n_mesh, int knn_mode,
                                float *dis, float *new_points, float *new_target,
                                float *jac_r, float *jac_t, float *pose_out, void *stream);

/* Pose hand-off between refinement stages (tools/train.py:141-145; the composition DenseFusion's evaluation
 * performs): pose_acc [F][12] (R row-major, t) <- pose_acc o pose_delta, i.e. R <- R Rd, t <- R td + t, where
 * pose_delta is the pose_out of the NEXT stage (estimated in the frame re-centred by pose_acc). */
int plr2_pose_compose(float *pose_acc, const float *pose_delta, int frames, void *stream);

/* ---- (6) evaluation metrics after the path ---------------------------------
 * ADD / ADD-S between two poses applied to the same model points, per frame
 * (src/estimation/errors.py:53-67 ADD, :69-84 ADDS; both poses are R|t [F][12], row-major R then t, the
 * layout plr2_pose_compose / pose_out produce; a point maps to R p + t as points.apply_transform does,
 * src/estimation/state.py:235-244):
 *   add[f]  = mean_j || (Ra p_j + ta) - (Rb p_j + tb) ||
 *   adds[f] = mean_j min_k || (Rb p_j + tb) - (Ra p_k + ta) ||     (model_1 = pose_a is the fitted set,
 *             model_2 = pose_b the queries, errors.py:80-83)
 * model_points [F][M][3] with model_stride floats between frames (0: one model shared by all frames).
 * add or adds may be NULL (not computed). FP32 arithmetic (the reference computes in float64).
 * workspace: plr2_pose_metric_workspace_bytes(frames) bytes, contents irrelevant. n_mesh <= 8192. */
size_t plr2_pose_metric_workspace_bytes(int frames);
/* Rotation and translation error between two poses per frame (src/estimation/errors.py:12-51):
 *   rot_err[f]   = || Rodrigues(Ra Rb^T) || = rotation angle of Ra Rb^T in [0, pi]   (errors.py:29-32)
 *   trans_err[f] = || ta - tb ||                                                     (errors.py:50)
 * computed as atan2(|axis vector|, trace - 1) (well conditioned at both ends). Either output may be NULL. */
int plr2_pose_errors(const float *pose_a, const float *pose_b, int frames, float *rot_err, float *trans_err,
                     void *stream);
int plr2_pose_metric(const float *model_points, long long model_stride, const float *pose_a,
                     const float *pose_b, int frames, int n_mesh, float *add, float *adds,
                     void *workspace, size_t workspace_bytes, void *stream);

/* Area under the accuracy-threshold curve of the YCB-Video toolbox
 * (replace_ycb_toolbox/plot_accuracy_keyframe.m:38-52 and VOCap :150-170): distances above max_distance
 * (0.1 there) count as misses, accuracy(d) = rank/n over the sorted distances, AUC = integral of the
 * accuracy envelope over [0, max_distance] divided by max_distance. Computed without a sort (rank and
 * predecessor of every distinct value by brute force over all pairs), accumulated in float64.
 *   out[0] = AUC in [0,1] (the script prints it x100), out[1] = fraction of distances < below (0.02 there).
 * n == 0 or no distance <= max_distance gives out[0] = 0. workspace: plr2_auc_workspace_bytes(n). */
size_t plr2_auc_workspace_bytes(int n);
int plr2_auc_f32(const float *dist, int n, double max_distance, double below, double *out,
                 void *workspace, size_t workspace_bytes, void *stream);

/* ---- (7) input step before the path ---------------------------------------
 * What datasets/ycb/dataset.py:177-197,215-219 (and src/loaders_v2/dataset_ycb.py:91-119) do per frame with
 * numpy on the host, for F frames on the device. Images are full frames [F][img_h][img_w]; bbox [F][4] int32 =
 * (rmin, rmax, cmin, cmax) from get_bbox (half-open, the crop is [rmin,rmax) x [cmin,cmax)).
 *
 * plr2_choose_from_mask: `choose = mask[rmin:rmax, cmin:cmax].flatten().nonzero()[0]` (ascending crop-flat
 *   indices), then exactly n_points of them:
 *     count <= n_points : np.pad(choose, (0, n_points - count), 'wrap')         (dataset.py:184)
 *     count >  n_points : an order-preserving subset of n_points entries          (dataset.py:179-182)
 *       keep != NULL: entry k of the nonzero list is kept iff keep[f*keep_stride + k] != 0 - the c_mask of the
 *                     reference, so a caller that draws it with numpy's RNG reproduces the reference exactly;
 *       keep == NULL: a uniformly random subset drawn on the device from (seed, frame, k) hash keys
 *                     (the n_points smallest keys; same distribution, not numpy's Mersenne-Twister stream).
 *   mask: uint8, non-zero = object pixel with valid depth (mask_depth * mask_label, dataset.py:126-128).
 *   count_out[f] = number of non-zero crop pixels (0: choose[f] is zero-filled; the reference raises there).
 *   workspace: plr2_choose_workspace_bytes(frames, max_crop_pixels) bytes.
 *
 * plr2_backproject: cloud[f][n] = ((col - cx) * z / fx, (row - cy) * z / fy, z), z = depth / cam_scale, with
 *   row = rmin + c / crop_w, col = cmin + c % crop_w for c = choose[f][n] (xmap holds the row index, ymap the
 *   column: dataset.py:83-84,186-194), every operation rounded to float32 exactly as numpy does on float32
 *   arrays; add_t [F][3] (float64, NULL = none) is added in float64 and rounded once (dataset.py:195-196,226).
 *   depth is uint16 (depth_is_u16 != 0, the PNG's type) or float32. cam [F][5] float32 = (cx, cy, fx, fy, scale).
 *
 * plr2_transform_target: target[f][j] = model_points[f][j] . R[f]^T + t[f] (+ add_t[f]) in float64, rounded
 *   to float32 (dataset.py:215-219,229); R [F][9] row-major, t [F][3], add_t [F][3] or NULL, all float64. *
 * plr2_mask_bbox: mask[f] = (label[f] == obj_id[f]) & (depth[f] != 0) (dataset.py:143-145, uint8 0/1), count[f] its
 *   number of set pixels (the `> minimum_num_pt` test of :146), bbox[f] = get_bbox(label[f] == obj_id[f])
 *   (dataset.py:251-289): tight box of the label pixels, each side snapped up to the next multiple of 40
 *   (border_list), centred, then shifted back inside the img_h x img_w image. No label pixel: bbox = 0,0,0,0
 *   (the reference raises IndexError). label is uint8 (the -label.png class ids). */
int plr2_mask_bbox(const void *depth, int depth_is_u16, const uint8_t *label, const int32_t *obj_id, int frames,
                   int img_h, int img_w, uint8_t *mask, int32_t *bbox, int32_t *count, void *stream);
size_t plr2_choose_workspace_bytes(int frames, int max_crop_pixels);
int plr2_choose_from_mask(const uint8_t *mask, int img_h, int img_w, const int32_t *bbox, int frames,
                          int n_points, const uint8_t *keep, long long keep_stride, unsigned long long seed,
                          int64_t *choose, int32_t *count_out, void *workspace, size_t workspace_bytes,
                          void *stream);
int plr2_backproject(const void *depth, int depth_is_u16, int img_h, int img_w, const int32_t *bbox,
                     const int64_t *choose, const float *cam, const double *add_t, int frames, int n_points,
                     float *cloud, void *stream);
int plr2_transform_target(const float *model_points, const double *R, const double *t, const double *add_t,
                          int frames, int n_mesh, float *target, void *stream);

/* ---- (2) dense contractions --------------------------------------------------------------
 * General FP32 GEMM with fused epilogue, C(m,n) (+)= f(sum_k A(m,k) B(k,n)):
 *   A(m,k) = A[m*sam + k*sak],  B(k,n) = B[k*sbk + n*sbn],  C(m,n) = C[m*ldc + n]
 *   f(x) = relu?( x + bias[n] + frame_bias[(m / rows_per_frame)*ldfb + n] ) * (relu_mask(m,n) > 0)
 * Replaces the Conv1d(k=1)/Linear library calls of lib/network.py:54-63,107-121,152-163,193-200
 * and their autograd (dX = dY W, dW = dY^T X).  split_k > 1 needs accumulate=1 (atomic adds). */
typedef struct plr2_gemm_desc {
  const float *A; long long sam, sak;
  const float *B; long long sbk, sbn;
  float *C; long long ldc;
  int M, N, K;
  const float *bias;                 /* [N] or NULL */
  const float *frame_bias;           /* [frames][ldfb] or NULL */
  int rows_per_frame; long long ldfb;
  const float *relu_mask; long long ldmask;   /* or NULL */
  int relu, accumulate, split_k;
} plr2_gemm_desc;
int plr2_gemm_f32(const plr2_gemm_desc *desc, void *stream);

/* Tensor-core contraction (tcgen05, sm_100a): C[M,N] (f32, ldc) = relu?(X W^T + bias) with both
 * operands in SPLIT-BF16: two bf16 planes hi = bf16(v), lo = bf16(v - hi), plane p at
 * base + p*plane_stride (elements), rows K-contiguous with pitch lda / ldw.  The product is evaluated
 * as X_hi W_hi + X_hi W_lo + X_lo W_hi with FP32 accumulation in TMEM.  K must be a multiple of 64. */
int plr2_split_bf16x2(const float *src, long long ld, long long rows, int cols, void *dst_planes,
                      long long ldd, long long plane_stride, void *stream);
int plr2_gemm_bf16x2(const void *A_planes, long long lda, long long a_plane, const void *W_planes,
                     long long ldw, long long w_plane, int M, int N, int K, const float *bias, int relu,
                     float *C, long long ldc, void *stream);
/* The weight-gradient form of the same contraction (autograd of the Conv1d(k=1)/Linear layers of
 * lib/network.py:54-63,107-121: dW = dY^T X, contraction over the points): C[M,N] (f32, ldc) = A^T B with
 * A = split-bf16 planes [k_rows][lda] (M columns used), B = planes [k_rows][ldb] (N columns used), read
 * MN-major by the tensor cores (no transposed copy).  k_rows is arbitrary; pitches multiples of 8 elements. */
int plr2_gemm_bf16x2_tn(const void *A_planes, long long lda, long long a_plane, const void *B_planes,
                        long long ldb, long long b_plane, int M, int N, int k_rows, float *C,
                        long long ldc, void *stream);

/* ---- (1)+(2) PoseNet / PoseRefineNet hot path ---------------------------------------------
 * Parameter tables: device pointers to the tensors of the reference's state dict, in this order
 * (weight then bias of each layer; Conv1d weights (out,in,1) and Linear weights (out,in) are both
 * row-major [out][in]):
 *   PoseNet       (lib/network.py:39-52,70-93): feat.conv1, feat.e_conv1, feat.conv2, feat.e_conv2,
 *                 feat.conv5, feat.conv6, conv1_r, conv1_t, conv1_c, conv2_r, conv2_t, conv2_c,
 *                 conv3_r, conv3_t, conv3_c, conv4_r, conv4_t, conv4_c              (18 layers)
 *   PoseRefineNet (lib/network.py:136-186): feat.conv1, feat.e_conv1, feat.conv2, feat.e_conv2,
 *                 feat.conv5, feat.conv6, conv1_r, conv1_t, conv2_r, conv2_t, conv3_r, conv3_t (12)
 * The same struct is used for gradients (float* written instead of read). */
#define PLR2_POSENET_LAYERS 18
#define PLR2_REFINER_LAYERS 12
typedef struct plr2_param_table {
  float *w[PLR2_POSENET_LAYERS];
  float *b[PLR2_POSENET_LAYERS];
} plr2_param_table;

#define PLR2_NET_POSENET 0
#define PLR2_NET_REFINER 1
/* precision of the large contractions */
#define PLR2_PREC_FP32 0      /* FP32 SIMT everywhere.  Forward and input gradients are bit-for-bit reproducible; the weight
                               * gradients are summed over the points by split-K partial products combined with atomicAdd,
                               * so their last bits depend on the arrival order (relative spread ~1e-7). */
#define PLR2_PREC_BF16X2 1    /* tcgen05 kind::f16 on operands split into two 16-bit planes (hi+lo, 3 MMAs), FP32 accumulate in
                               * TMEM.  Inference keeps activations and packed weights as IEEE fp16 planes (22 mantissa bits;
                               * environment PLR2_INFER_FMT=bf16 selects bf16 planes, 16 bits - set it before the first
                               * plr2_net_pack, the planes of the packed buffer must match).  With training != 0 activations stay FP32:
                               * the forward runs on the FP32 kernel (its ReLU signs are the masks of the backward),
                               * the operands of every input-gradient / weight-gradient contraction are packed into
                               * scratch planes of the workspace (n_points >= 128, else FP32 SIMT). */
#define PLR2_PREC_BF16X2_ALL 2 /* as above, and the training forward on the tensor cores too; pre-activations inside the
                               * operand-error bound (2^-14 |x||w|, ~1e-3 of them) are recomputed in FP32 before the ReLU
                               * so that the masks of the backward agree with an FP32 forward */

/* Packed weight buffer (floats) and workspace (bytes) sizes. */
size_t plr2_net_packed_floats(int net, int num_obj);
size_t plr2_net_workspace_bytes(int net, int frames, int n_points, int num_obj, int training);
/* Gather the state-dict tensors into the packed layout: packed[0, T) holds the FP32 copy and, when
 * split_bf16 != 0, packed[T, 2T) (T = plr2_net_packed_floats/2) the hi/lo 16-bit planes of the same layout (fp16 by
 * default, see PLR2_PREC_BF16X2; they feed the inference path only - training packs its operands per contraction). */
int plr2_net_pack(int net, const plr2_param_table *params, int num_obj, float *packed, int split_bf16,
                  void *stream);
/* Scatter a packed gradient buffer back into per-parameter tensors (overwrites them). */
int plr2_net_unpack_grads(int net, const float *packed_grad, const plr2_param_table *grads, int num_obj,
                          void *stream);
/* The same scatter ADDED onto the tensors (autograd's AccumulateGrad for all parameters of a network in one launch
 * instead of one element-wise add per parameter: tools/train.py:150-155 accumulates 8 frames before a step). */
int plr2_net_accumulate_grads(int net, const float *packed_grad, const plr2_param_table *grads, int num_obj,
                              void *stream);

/* Input-range status of the LAST plr2_posenet_forward / plr2_refiner_forward that used `workspace`.  torch.gather
 * (lib/network.py:102) and index_select (lib/network.py:123-126, 202-204) raise on an out-of-range index; a kernel
 * cannot, so the forward clamps the index (no fault), sets a flag in the first word of its workspace, and this call
 * copies the flags to the host: it SYNCHRONISES `stream` (debug / validation use, not for the hot loop).
 * *flags == 0: every choose index was in [0, hw) and every class index in [0, num_obj). */
#define PLR2_STATUS_CHOOSE_RANGE 1u
#define PLR2_STATUS_OBJ_RANGE 2u
int plr2_net_input_status(const void *workspace, unsigned int *flags, void *stream);

/* PoseNet.forward without the backbone (lib/network.py:98-132) for `frames` frames:
 *   out_img [F][32][hw] f32 backbone output; choose [F][N] i64; x [F][N][3]; obj [F] i64
 *   -> pred_r [F][N][4], pred_t [F][N][3], pred_c [F][N][1], emb [F][32][N]
 * training != 0 keeps every activation in the workspace for plr2_posenet_backward (pass the same
 * `precision` to both: the backward runs its contractions the way the forward did). */
int plr2_posenet_forward(const float *packed, const float *out_img, const int64_t *choose, const float *x,
                         const int64_t *obj, int frames, int n_points, int hw, int num_obj,
                         void *workspace, size_t workspace_bytes, float *pred_r, float *pred_t,
                         float *pred_c, float *emb, int training, int precision, void *stream);
/* Backward: upstream grads of pred_r/t/c -> packed parameter grads (overwritten) and, if
 * g_out_img != NULL, the gradient w.r.t. the backbone output (must be zero-filled by the caller;
 * duplicates in `choose` accumulate).  Uses the workspace left by the training forward. */
int plr2_posenet_backward(const float *packed, const int64_t *choose, const float *x, const int64_t *obj,
                          int frames, int n_points, int hw, int num_obj, void *workspace,
                          size_t workspace_bytes, const float *g_pred_r, const float *g_pred_t,
                          const float *g_pred_c, float *packed_grad, float *g_out_img, int precision,
                          void *stream);

/* Overlap of the gradient all-reduce with the backward pass ("bucketed so head-layer buckets overlap the
 * feature-layer backward"). plr2_net_grad_regions gives the packed-gradient range [early_begin, early_end) (floats)
 * holding the head-layer weights - 83 % of PoseNet's 16 MB - which plr2_posenet_backward completes first.
 * plr2_set_backward_stage_event arms a cudaEvent_t (one-shot, per calling thread) that the NEXT
 * plr2_posenet_backward records on its stream once that range is final; a side stream that waits on it can
 * all-reduce the range while the trunk backward still runs (plr2_b200/ops.py: PoseNetFn.backward). */
int plr2_set_backward_stage_event(void *cuda_event);
int plr2_net_grad_regions(int net, int num_obj, size_t *early_begin, size_t *early_end);

/* PoseRefineNet.forward (lib/network.py:187-206): x [F][N][3], emb [F][32][N], obj [F]
 *   -> pred_r [F][4], pred_t [F][3]
 * training: 0 inference, 1 training (activations kept for plr2_refiner_backward), or
 * PLR2_REFINER_NEXT_ITER: inference for refine iteration >= 2 (tools/train.py:141-145 feeds the SAME emb to
 * every iteration): the caller guarantees that emb, the packed parameters, frames/n_points and the workspace
 * are those of the immediately preceding plr2_refiner_forward call on this workspace, so the emb branch
 * (e_conv1, e_conv2) still held there is reused instead of recomputed. Ignored (full forward) on the FP32 path. */
#define PLR2_REFINER_NEXT_ITER 2
int plr2_refiner_forward(const float *packed, const float *x, const float *emb, const int64_t *obj,
                         int frames, int n_points, int num_obj, void *workspace, size_t workspace_bytes,
                         float *pred_r, float *pred_t, int training, int precision, void *stream);
int plr2_refiner_backward(const float *packed, const float *x, const float *emb, const int64_t *obj,
                          int frames, int n_points, int num_obj, void *workspace, size_t workspace_bytes,
                          const float *g_pred_r, const float *g_pred_t, float *packed_grad, int precision,
                          void *stream);

/* ---- batched pose Kalman filter ----------------------------------------------------------
 * The per-sequence loop of src/helper/postprocess.py:35-61 over Kalman_Filter.update / .predict
 * (src/estimation/filter.py:132-189), State_R3xQuat (src/estimation/state.py:100-185) and Linear_Motion
 * (src/estimation/motion.py:15-60) for `sequences` independent sequences of `steps` observations, float64 like the
 * reference.  State = t xyz | q xyzw.
 *   prior, prior_var [S][7]         set_prior(prior, covariance) with a 7-vector covariance (filter.py:148-157)
 *   obs, sensor_var  [S][T][7]      update(observation, covariance_sensor); kf_sequence passes var_sensor / obs_c
 *   motion_var       [7]            variance_motion (diagonal of Q)
 *   lengths          [S] or NULL    observations of sequence s used (<= steps); outputs beyond stay untouched
 *   a_t, a_r                        Linear_Motion params (motion.py:18-26)
 *   filtered  [S][T][7]             the state update() returns for each observation
 *   p_diag    [S][T][7] or NULL     diagonal of the covariance after each update (the reference's element-wise update,
 *                                   filter.py:186, leaves nothing else)
 *   predicted [S][T][7] or NULL     the state predict() appends after each observation
 * Deviation (stated): the covariance is propagated with the ANALYTIC Jacobian of the pose composition; the reference
 * differences it with a step of machine epsilon (state.py:150-161), which returns rounding noise. */
int plr2_kalman_filter_f64(const double *prior, const double *prior_var, const double *obs,
                           const double *sensor_var, const double *motion_var, const int32_t *lengths,
                           double a_t, double a_r, int sequences, int steps,
                           double *filtered, double *p_diag, double *predicted, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PLR2_B200_H */
