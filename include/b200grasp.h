/*
 * b200grasp.h -- C ABI of libb200grasp.so: the B200 (sm_100a) implementation of
 * GraspNet-baseline's point-cloud geometry hot path.
 *
 * Conventions (all entry points):
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless
 *     the comment says HOST; tensors are contiguous, row-major, fp32 / int32
 *     unless stated;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default
 *     stream); launches are asynchronous, nothing synchronises, nothing
 *     allocates (scratch is caller-provided), nothing calls exit();
 *   - return 0 on success, a positive cudaError_t value if a launch failed, or
 *     a negative B200_E* code for bad arguments.  b200_error_string() renders
 *     either.  (The reference prints and exit(-1)s on a launch failure,
 *     pointnet2/_ext_src/include/cuda_utils.h:35-44; knn prints and carries on,
 *     knn/src/knn.h:40-46.  Here the host wrapper raises RuntimeError.)
 *   - outputs the reference zero-fills (torch::zeros) are fully written by the
 *     callee, so the caller may pass uninitialised memory.
 *
 * Each entry point names the reference interface it replaces (paths relative
 * to the reference repository root).
 */
#ifndef B200GRASP_H_
#define B200GRASP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_EINVAL (-1)      /* bad size / null pointer                       */
#define B200_ECAPACITY (-2)   /* problem larger than this build supports       */
#define B200_EWORKSPACE (-3)  /* workspace too small                           */
#define B200_ESTATIC (-4)     /* larger than a *_static entry point's built-in scratch */

int b200_version(void);
const char *b200_error_string(int code);
/* Number of kernels the library has launched so far in this process (diagnostic). */
unsigned long long b200_launch_count(void);

/* Replaces furthest_point_sampling(points[B,N,3], nsamples) -> [B,nsamples] i32
 * (pointnet2/_ext_src/src/sampling.cpp:70-91, include/sampling.h:11; kernel
 * sampling_gpu.cu:74-178).  First index 0; points with |p|^2 <= 1e-3 never
 * selected; ties resolved exactly as the legacy block reduction does.
 * The legacy (b,n) `temp` buffer is not needed (min-distances live on chip). */
int b200_furthest_point_sampling(const float *xyz, int b, int n, int m, int32_t *idx, void *stream);

/* Replaces gather_points(points[B,C,N], idx[B,M]) -> [B,C,M]
 * (sampling.cpp:20-43, sampling.h:9; sampling_gpu.cu:13-25). */
int b200_gather_points(const float *points, const int32_t *idx, int b, int c, int n, int m,
                       float *out, void *stream);

/* Replaces gather_points_grad(grad_out[B,C,M], idx[B,M], n) -> [B,C,n]
 * (sampling.cpp:45-69, sampling.h:10; sampling_gpu.cu:39-52).  grad_points is
 * zeroed by the callee, then scatter-added. */
int b200_gather_points_grad(const float *grad_out, const int32_t *idx, int b, int c, int n, int m,
                            float *grad_points, void *stream);

/* Speculative FPS (PointNet++ feeds SA2..SA4 the points SA1 chose IN THE ORDER it chose them, models/backbone.py:
 * 110-119; FPS over such a prefix returns 0,1,2,... unless a tie or the skip rule intervenes).
 *  b200_fps_prefix_check: decides, without the serial chain, whether FPS over the first n points of every cloud
 *    (clouds `cloud_stride` >= n points apart, so a prefix VIEW of a larger array works) returning m samples is the
 *    identity: verdict[c*m] = 0 if so, -1 if the chain must run (verdict is b*m int32 of scratch).
 *  b200_furthest_point_sampling_spec: b200_furthest_point_sampling that takes the shortcut for cloud c iff every
 *    given verdict array (up to three, entry [c*stride]; NULL = unused, from slot 0) says 0.  Several arrays are
 *    for levels whose input is a prefix only if the levels above returned the identity as well; this lets all the
 *    checks of a backbone run concurrently right after SA1.  The result is always the exact FPS answer. */
int b200_fps_prefix_check(const float *xyz, int b, int n, int m, long long cloud_stride, int32_t *verdict,
                          void *stream);
int b200_furthest_point_sampling_spec(const float *xyz, int b, int n, int m, const int32_t *verdict0, int stride0,
                                      const int32_t *verdict1, int stride1, const int32_t *verdict2, int stride2,
                                      int32_t *idx, void *stream);

/* SA-module fast path: new_xyz[B,M,3] = xyz[B,N,3] rows picked by idx[B,M], and (if out_t != NULL) the same
 * points channel-major [B,3,M] -- what gather_operation(xyz^T, inds) and its .transpose(1,2).contiguous() give
 * (pointnet2_modules.py:230-240), in one launch. */
int b200_gather_xyz(const float *xyz, const int32_t *idx, int b, int n, int m, float *out, float *out_t,
                    void *stream);

/* ---- group -> shared MLP -> max-pool as one kernel (csrc/mlp_fused.cu) -------------------
 * Replaces, in eval mode and for modules whose grouped input is xyz only (SA1, CloudCrop), the chain
 * QueryAndGroup / CylinderQueryAndGroup.forward (pointnet2_utils.py:313-372, 488-555) -> pt_utils.SharedMLP
 * (pytorch_utils.py:11-36) -> F.max_pool2d over nsample (pointnet2_modules.py:243-257, models/modules.py:127-141).
 *   xyz[B,N,3], new_xyz[B,M,3], rot[B,M,9] or NULL (rotate_xyz), idx[G,B,M,ns] (G = 1, or CloudCrop's depth lists)
 *   radius > 0: grouped_xyz /= radius (normalize_xyz); <= 0: left as is
 *   layer 1: w1[C1,3], b1[C1]; layers 2, 3: weights packed by b200_mlp_pack_weights, b2[C2], b3[C3]; BatchNorm is
 *   folded into (w, b) by the caller.  out[B,C3,M,G] (= [B,C3,M] for G = 1; CloudCrop's view(B,-1,num_seed,num_depth)).
 *   out_t (or NULL): the same features channels-last, [B,M,G,C3] -- the layout b200_group_feat_mlp_maxpool reads.
 *   passes = 3: fp32-grade ("3xTF32" on the tensor cores), passes = 1: TF32 (cuDNN's default for the reference convs).
 * Supported: ns in {16,32,64,128}; C1, C2, C3 multiples of 16, all <= 256 (else B200_ECAPACITY). */
int b200_group_mlp_maxpool(const float *xyz, const float *new_xyz, const float *rot, const int32_t *idx, int g, int b,
                           int n, int m, int ns, float radius, const float *w1, const float *b1,
                           const float *w2_packed, const float *b2, const float *w3_packed, const float *b3, int c1,
                           int c2, int c3, int passes, float *out, float *out_t, void *stream);
/* The same for modules whose grouped input is [xyz | point features] (SA2-SA4: QueryAndGroup with features,
 * pointnet2_utils.py:352-358 torch.cat([grouped_xyz, grouped_features], dim=1)); all three layers on the tensor cores.
 *   feat_t[B,N,C]: the features CHANNELS-LAST (a row's C values contiguous: one neighbour = C/4 16-byte loads;
 *   the reference layout [B,C,N] is transposed by the caller, or taken from out_t of the previous module);
 *   w1_packed: b200_mlp_pack_weights of the first conv's weight with its input columns reordered to
 *   [features (C) | x y z | 13 zero columns] (cin = C + 16).  C a multiple of 16, C1 + C2 + C3 <= 512. */
int b200_group_feat_mlp_maxpool(const float *xyz, const float *new_xyz, const float *feat_t, const int32_t *idx, int b,
                                int n, int m, int ns, int c, float radius, const float *w1_packed, const float *b1,
                                const float *w2_packed, const float *b2, const float *w3_packed, const float *b3,
                                int c1, int c2, int c3, int passes, float *out, float *out_t, void *stream);
/* w[cout,cin] fp32 (conv weight, BN folded) -> packed[2*cout*cin] in the order the kernel's bulk copies read:
 * K-slabs of 16 input channels, each [hi | lo][k-chunk of 4][cout][4], hi = tf32(w), lo = tf32(w - hi).
 * cout a multiple of 16 up to 256, cin a multiple of 16 up to 512. */
int b200_mlp_pack_weights(const float *w, int cout, int cin, float *packed, void *stream);

/* ---- The same modules, and the stacks of 1x1 convolutions between them, in "fp16x3" arithmetic (csrc/mlp_f16.cu):
 * operands split hi + lo in fp16 with exact power-of-two scales (weights per layer, activations per warp of 32 rows
 * from a rigorous bound), three tcgen05.mma kind::f16 products per K = 16 step, fp32 accumulation -- fp32-grade like
 * 3xTF32 at twice the tensor rate and half the accumulation steps; weights stay resident in shared memory when they
 * fit.  One tensor-core layer y = act(W x + b), BatchNorm folded by the caller: */
typedef struct {
  const void *w_packed;       /* b200_mlp_f16_pack of W (cout, cin) zero-padded to (n_pad, k_pad)                  */
  const float *bias;          /* [n_pad], zero padded                                                              */
  int k_pad, n_pad;           /* k_pad % 32 == 0 (= the previous layer's n_pad rounded up to 32), n_pad % 16 == 0,
                                 16 <= n_pad <= 256                                                                */
  int cout;                   /* real output channels                                                              */
  int relu;                   /* apply ReLU                                                                        */
  int w_exp;                  /* W was multiplied by 2^w_exp before the split (so that max |W| lands near 2^13)    */
  float w_norm1, b_max;       /* max_c sum_k |W[c,k]| and max_c |b[c]| (of the unscaled W, b): the range bound     */
  float *y;                   /* pointwise stacks: this layer's output [B, *, n] (NULL = not written; the last
                                 layer's must be set); grouped modules: NULL                                       */
  long long y_batch_stride;   /* floats between batches of y                                                       */
} b200_mlp_layer;
size_t b200_mlp_f16_packed_bytes(int n_pad, int k_pad);
/* w[cout, cin] (row stride ldw) -> packed.  feat_channels = C > 0: the input columns are [x y z | C features] (the
 * reference's concatenation order, pointnet2_utils.py:351-355) and are packed as [features | x y z | pad]. */
int b200_mlp_f16_pack(const float *w, int cout, int cin, int ldw, int n_pad, int k_pad, int w_exp, int feat_channels,
                      void *packed, void *stream);
/* b200_group_mlp_maxpool in fp16x3: layer 1 (w1[c1,3], b1[c1], on the CUDA cores; w1_norm1 = max_c (|w|_1 of row c),
 * b1_max = max |b1|), then layers[0..1] on the tensor cores (layers[0].k_pad = c1 rounded up to 32).
 * out_row_bound (may be NULL): [B * m * g][8] f32, 32-byte aligned -- the ROW-BOUND RIDER of out_t: for every output
 * row eight partial maxima of |y| over its channels (one per writing warp, unused ones 0; the row's bound is their
 * maximum).  Handed to the next module as feat_row_bound it replaces that module's pass over each gathered row for
 * the row's magnitude (the fp16 scale is derived from it): same value, so the same bits come out. */
int b200_group_mlp_maxpool_f16(const float *xyz, const float *new_xyz, const float *rot, const int32_t *idx, int g,
                               int b, int n, int m, int ns, float radius, const float *w1, const float *b1, int c1,
                               float w1_norm1, float b1_max, const b200_mlp_layer *layers, float *out, float *out_t,
                               float *out_row_bound, void *stream);
/* b200_group_feat_mlp_maxpool in fp16x3: layers[0..2], layers[0] packed with feat_channels = c; c % 16 == 0 and
 * feat_t 32-byte aligned (the rows are gathered with 256-bit loads).  feat_row_bound (may be NULL): the rider written
 * beside feat_t by the call that produced it ([B * n][8]); out_row_bound as above. */
int b200_group_feat_mlp_maxpool_f16(const float *xyz, const float *new_xyz, const float *feat_t,
                                    const float *feat_row_bound, const int32_t *idx, int b, int n, int m, int ns, int c,
                                    float radius, const b200_mlp_layer *layers, float *out, float *out_t,
                                    float *out_row_bound, void *stream);
/* Replaces the Conv1d / Conv2d(1x1) [+ BatchNorm + ReLU] stacks over point columns between the geometry ops
 * (pointnet2/pointnet2_modules.py:409-414, models/modules.py:52-57, 170-177, 215-222): x = [x1[B,c1,n] | x2[B,c2,n]]
 * through 1..3 layers in one launch, activations resident on the SM; each layer with y != NULL is written
 * [B, cout, n] (channel-major, like the input).
 * nheads = 1 or 2: two stacks of the SAME padded shapes (own weights, biases, outputs) over the same input run as one
 * launch -- layers[0 .. nlayers) is head 0, layers[nlayers .. 2 nlayers) head 1 (OperationNet | ToleranceNet over the
 * crop features; the two halves of a layer wider than 256 channels). */
int b200_pointwise_stack_f16(const float *x1, int c1, const float *x2, int c2, int b, int n,
                             const b200_mlp_layer *layers, int nlayers, int nheads, void *stream);
/* Scheduling knob of the fp16x3 entry points (not part of the reference's interface): at least k tiles per CTA for the
 * launches with features / dense inputs, 4 k for the xyz-only modules when k > 1 (default k = 1).  A launch of few tiles
 * (SA3 / SA4 / the heads of ONE scene) then uses fewer CTAs for longer: less SM-time per scene for a stream of scenes, a
 * longer launch for a scene alone.  Results do not depend on it.
 * Process-wide, applies to the launches (or graph captures) that follow. */
void b200_mlp_f16_set_tiles_per_cta(int k);

/* Replaces ball_query(new_xyz[B,M,3], xyz[B,N,3], radius, nsample) -> [B,M,nsample] i32
 * (ball_query.cpp:13-37, ball_query.h:9-10; ball_query_gpu.cu:14-49). */
int b200_ball_query(const float *new_xyz, const float *xyz, int b, int n, int m, float radius,
                    int nsample, int32_t *idx, void *stream);

/* Replaces cylinder_query(new_xyz[B,M,3], xyz[B,N,3], rot[B,M,9], radius, hmin, hmax, nsample)
 * -> [B,M,nsample] i32  (cylinder_query.cpp:10-37, cylinder_query.h:6-7;
 * cylinder_query_gpu.cu:9-57). */
int b200_cylinder_query(const float *new_xyz, const float *xyz, const float *rot, int b, int n,
                        int m, float radius, float hmin, float hmax, int nsample, int32_t *idx,
                        void *stream);

/* CloudCrop fast path (models/modules.py:127-130 calls cylinder_query once per
 * hmax with otherwise identical inputs): one scan emits all `nh` (<= 8) lists.
 * hmax is a HOST array of nh floats; idx is [nh,B,M,nsample]; slice h equals
 * b200_cylinder_query(..., hmax[h], ...). */
int b200_cylinder_query_multi(const float *new_xyz, const float *xyz, const float *rot, int b,
                              int n, int m, float radius, float hmin, const float *hmax, int nh,
                              int nsample, int32_t *idx, void *stream);

/* Replaces pred_decode(end_points) (models/graspnet.py:77-134, with batch_viewpoint_params_to_matrix,
 * utils/loss_utils.py:68-96) for a batch of b scenes in one launch:
 *   objectness [b,2,ns]; score / angle_cls / width / tolerance [b,A,ns,D] (A = num_angle, D = num_depth <= 8);
 *   xyz (fp2_xyz) and view_xyz (grasp_top_view_xyz) [b,ns,3]; all fp32.
 *   rows [b,ns,17] OUT: the rows of the seeds whose objectness argmax is 1, in seed order, as
 *   [score, width, 0.02, depth, R(9 row-major), centre(3), -1]; nrows [b] i32 OUT: how many per scene
 *   (rows beyond it are not written).  No host synchronisation. */
int b200_pred_decode(const float *objectness, const float *score, const float *angle_cls, const float *width,
                     const float *tolerance, const float *xyz, const float *view_xyz, int b, int ns,
                     int num_angle, int num_depth, float *rows, int32_t *nrows, void *stream);

/* 1x1 convolution over point columns with the bias (eval-mode BatchNorm folded in) and optional ReLU in the epilogue:
 *   y[b,co,j] = act(bias[co] + sum_ci w[co,ci] * x[b,ci,j]),  x = the channel concatenation [x1 (c1 rows) | x2 (c2 rows)]
 * (x2 may be NULL with c2 = 0).  One launch per layer of PointnetFPModule's SharedMLP (pointnet2_modules.py:403-414,
 * pytorch_utils.py:11-36), ApproachNet (models/modules.py:52-57), OperationNet / ToleranceNet (:170-177, :215-222).
 * w [cout, ldw] with ldw >= c1+c2 the row stride (columns c1+c2..ldw-1 must be ZERO: rows padded to a multiple of 4
 * floats take the 16-byte cp.async path), bias [cout], x1 [b,c1,n], x2 [b,c2,n], y [b,cout,n]; fp32, exact FMAs. */
int b200_pointwise_conv(const float *w, int ldw, const float *bias, const float *x1, int c1, const float *x2, int c2,
                        int b, int n, int cout, int relu, float *y, void *stream);

/* ApproachNet's tail (models/modules.py:60-73): features [b,2+v,s] fp32 (conv3's output) -> per seed the first
 * maximum over the v view scores: top_inds [b,s] i64, top_score [b,s], vp_xyz [b,s,3] = views[top], vp_rot [b,s,9] =
 * frames[top].  views [v,3] = generate_grasp_views(v) (utils/loss_utils.py:41-66) and frames [v,9] =
 * batch_viewpoint_params_to_matrix(-views, 0) (:68-96) are caller-provided device tables (computed once). */
int b200_view_select(const float *features, const float *views, const float *frames, int b, int s, int v,
                     long long *top_inds, float *top_score, float *vp_xyz, float *vp_rot, void *stream);

/* PointnetFPModule's interpolation weights (pointnet2/pointnet2_modules.py:394-397 after pointnet2_utils.py:140):
 * dist2 [n,3] squared distances from b200_three_nn -> dist [n,3] = sqrt (may be NULL), weight [n,3] =
 * (1/(dist+1e-8)) / sum over the three. */
int b200_three_nn_weights(const float *dist2, int n, float *dist, float *weight, void *stream);

/* Device-resident post-processing: pred_decode's rows -> collision mask -> surviving rows -> NMS -> kept rows, with
 * every row count read from / written to device memory (no host synchronisation between the stages).
 *  b200_collision_mask_device: ModelFreeCollisionDetector.detect's default return (collision_detector.py:37-112) for
 *    fp32 grasp rows [m_max,17] of which the first *m_dev are valid; mask[g] = 1 iff global_iou > collision_thresh,
 *    with the reference's fp32 threshold / volume arithmetic.  scene as in b200_collision_counts_dn.
 *    workspace: b200_collision_mask_workspace(m_max) bytes.
 *  b200_compact_rows: out = rows[~drop] in order (test.py:96), *nout = how many.
 *  b200_grasp_nms_device: b200_grasp_nms with the row count on the device; workspace b200_grasp_nms_workspace(m_max).
 *  b200_gather_rows: out[i] = rows[index[i]] for i < *k_dev (the GraspGroup that nms() returns). */
size_t b200_collision_mask_workspace(int m_max);
int b200_collision_mask_device(const double *scene, const int32_t *n_dev, int n_max, const float *grasps,
                               const int32_t *m_dev, int m_max, double voxel_size, double approach_dist,
                               double collision_thresh, uint8_t *mask, void *workspace, size_t workspace_bytes,
                               void *stream);
int b200_compact_rows(const float *rows, int cols, const uint8_t *drop, const int32_t *m_dev, int m_max, float *out,
                      int32_t *nout, void *stream);
int b200_grasp_nms_device(const float *grasps, const int32_t *m_dev, int m_max, double translation_thresh,
                          double cos_rotation_thresh, int32_t *keep, int32_t *nkeep, void *workspace,
                          size_t workspace_bytes, void *stream);
int b200_gather_rows(const float *rows, int cols, const int32_t *index, const int32_t *k_dev, int k_max, float *out,
                     void *stream);

/* ---- training-path label generation fused with the nearest-grasp-point search (csrc/labels.cu).
 * Replaces the index_select / cat / gather chain of utils/label_generation.py:58-118 (process_grasp_labels) and
 * :125-149 (match_grasp_view_and_label) for everything the loss reads, without building the (B,Ns,V,A,D[,3]) tensors.
 * A batch is described by device tables: an object table (one entry per object of every scene, filled on the HOST
 * with b200_label_object_table_set and copied to the device: labels[Np,V,A,D], offsets[Np,V,A,D,3], tolerance[Np,V,A,D]
 * device pointers, view_inds[V] i32 = for every template view of the scene frame the nearest rotated template view
 * (label_generation.py:58-60, knn k = 1), rot9 = device pointer to the pose rotation), obj_first[B] (first object of scene b), the merged grasp
 * points of all scenes (point_obj[ ] = object within its scene, point_local[ ] = row within the object,
 * point_first[B]) and nn[B,Ns] i64 = the seed's nearest merged point within its scene (knn k = 1, 0-based, :84-86).
 *  b200_label_view_stats: lmin[B,Ns,V] = smallest label of the row (nn, view_inds[v]) with label > 0 and width <=
 *    max_width (+inf if none); *umax_key = the largest gathered label of the batch (order-preserving integer image;
 *    zeroed by the callee first).
 *  b200_label_view_scores: scores[n] = log(u_max / lmin) or 0 -- 'batch_grasp_view_label' (:113-118); *umax_out = u_max.
 *  b200_label_top_gather: for top_view[B,Ns] i64: out_label[B,Ns,A*D] (transformed), out_offset[B,Ns,A*D,3],
 *    out_tol[B,Ns,A*D], out_rot[B,Ns,9] = pose rotation x view_rot_tmpl[view_inds[top]] (view_rot_tmpl[V,9] = the
 *    template frames, loss_utils.py:68-96). */
size_t b200_label_object_table_bytes(int n_objects);
int b200_label_object_table_set(void *host_table, int i, const float *labels, const float *offsets,
                                const float *tolerance, const int32_t *view_inds, const float *rot9);
int b200_label_view_stats(const void *objs, const int32_t *obj_first, const int32_t *point_obj,
                          const int32_t *point_local, const int32_t *point_first, const long long *nn, int b, int ns,
                          int v, int ad, float max_width, float *lmin, uint32_t *umax_key, void *stream);
int b200_label_view_scores(const float *lmin, const uint32_t *umax_key, long long n, float *scores, float *umax_out,
                           void *stream);
int b200_label_top_gather(const void *objs, const int32_t *obj_first, const int32_t *point_obj,
                          const int32_t *point_local, const int32_t *point_first, const long long *nn,
                          const long long *top_view, const float *view_rot_tmpl, const uint32_t *umax_key, int b, int ns,
                          int v, int ad, float max_width, float *out_label, float *out_offset, float *out_tol,
                          float *out_rot, void *stream);

/* Grid-accelerated forms of b200_ball_query / b200_cylinder_query_multi: identical results (the first
 * nsample hits IN INDEX ORDER, padded with the first hit, zeros if none), but the cloud is first binned
 * into a uniform grid held in `workspace` (b200_query_grid_workspace(b, n) bytes of device memory, no
 * alignment beyond cudaMalloc's) and each centre only visits the cells its bounding sphere touches.
 * Inputs the grid cannot prune (non-finite radius, clouds of >= 2^24 points) and centres whose `rot` is
 * not orthonormal are answered by the linear scan, so the result never depends on which form ran.
 *   *_grid      : bin (cell edge = half the query's bounding radius), then query.
 *   *_prebuilt  : query a grid that b200_query_grid_build made of the SAME xyz (b, n) earlier on the same
 *                 stream order; any cell edge gives the same lists, only the speed differs.  This is how one
 *                 binning serves SA1's ball query and CloudCrop's cylinder query (both read the full cloud,
 *                 models/backbone.py:93, models/modules.py:127-130). */
size_t b200_query_grid_workspace(int b, int n);
int b200_query_grid_build(const float *xyz, int b, int n, float cell, void *workspace, size_t workspace_bytes,
                          void *stream);
int b200_ball_query_grid(const float *new_xyz, const float *xyz, int b, int n, int m, float radius,
                         int nsample, int32_t *idx, void *workspace, size_t workspace_bytes, void *stream);
int b200_ball_query_prebuilt(const float *new_xyz, const float *xyz, int b, int n, int m, float radius,
                             int nsample, int32_t *idx, const void *grid, size_t grid_bytes, void *stream);
int b200_cylinder_query_multi_grid(const float *new_xyz, const float *xyz, const float *rot, int b, int n,
                                   int m, float radius, float hmin, const float *hmax, int nh, int nsample,
                                   int32_t *idx, void *workspace, size_t workspace_bytes, void *stream);
int b200_cylinder_query_multi_prebuilt(const float *new_xyz, const float *xyz, const float *rot, int b, int n,
                                       int m, float radius, float hmin, const float *hmax, int nh,
                                       int nsample, int32_t *idx, const void *grid, size_t grid_bytes,
                                       void *stream);

/* Replaces group_points(points[B,C,N], idx[B,M,ns]) -> [B,C,M,ns]
 * (group_points.cpp:17-40, group_points.h:9; group_points_gpu.cu:13-33). */
int b200_group_points(const float *points, const int32_t *idx, int b, int c, int n, int m, int ns,
                      float *out, void *stream);

/* Replaces group_points_grad(grad_out[B,C,M,ns], idx[B,M,ns], n) -> [B,C,n]
 * (group_points.cpp:42-65, group_points.h:10; group_points_gpu.cu:48-69). */
int b200_group_points_grad(const float *grad_out, const int32_t *idx, int b, int c, int n, int m,
                           int ns, float *grad_points, void *stream);

/* Replaces three_nn(unknowns[B,n,3], knows[B,m,3]) -> {dist2[B,n,3] f32, idx[B,n,3] i32}
 * (interpolate.cpp:19-45, interpolate.h:11; interpolate_gpu.cu:14-64).  dist2 is
 * the SQUARED distance; the host wrapper applies sqrt (pointnet2_utils.py:140). */
int b200_three_nn(const float *unknown, const float *known, int b, int n, int m, float *dist2,
                  int32_t *idx, void *stream);

/* Replaces three_interpolate(points[B,c,m], idx[B,n,3], weight[B,n,3]) -> [B,c,n]
 * (interpolate.cpp:47-75, interpolate.h:12-13; interpolate_gpu.cu:77-106). */
int b200_three_interpolate(const float *points, const int32_t *idx, const float *weight, int b,
                           int c, int m, int n, float *out, void *stream);

/* Replaces three_interpolate_grad(grad_out[B,c,n], idx, weight, m) -> [B,c,m]
 * (interpolate.cpp:76-104, interpolate.h:14-15; interpolate_gpu.cu:121-148). */
int b200_three_interpolate_grad(const float *grad_out, const int32_t *idx, const float *weight,
                                int b, int c, int n, int m, float *grad_points, void *stream);

/* Replaces knn(ref[B,D,Nr], query[B,D,Nq], idx[B,k,Nq] i64 OUT) (knn/src/knn.h:12-65,
 * knn/src/cuda/knn.cu:217-263).  1-based indices, stable top-k (ties -> lower
 * reference index).  No Nr x Nq scratch matrix.  k <= 32; rows k' >= Nr are
 * left untouched as in the reference. */
int b200_knn(const float *ref, const float *query, int b, int dim, int nr, int nq, int k,
             int64_t *ind, void *stream);

/* Integer core of ModelFreeCollisionDetector.detect (utils/collision_detector.py:69-100,119):
 * scene[N,3] f64, T[M,3] f64, R[M,9] f64 (row-major), thr[M,10] f64 = the ten
 * half-space bounds the host derives with the reference's dtype rules
 * (-h/2, h/2, d-fl, d, -(w/2+fw), -w/2, w/2+fw, w/2, d-fl-fw, d-fl-fw-ad).
 * counts[M,6] i32 = #points in {global, left, right, bottom, shifting, inner}. */
int b200_collision_counts(const double *scene, int n, const double *T, const double *R,
                          const double *thr, int m, int32_t *counts, void *stream);

/* Same, with the scene length read from DEVICE memory (*n_dev <= n_max) so the call can follow
 * b200_voxel_down_sample without a host round trip. */
int b200_collision_counts_dn(const double *scene, const int32_t *n_dev, int n_max, const double *T,
                             const double *R, const double *thr, int m, int32_t *counts, void *stream);

/* Same, with the per-grasp inputs in ONE row-major [M,22] fp64 buffer: T (3) | R (9) | thr (10) -- the layout
 * ModelFreeCollisionDetector.detect uploads with a single copy.  n_dev may be NULL (then n_max is the length). */
int b200_collision_counts_packed(const double *scene, const int32_t *n_dev, int n_max, const double *packed,
                                 int m, int32_t *counts, void *stream);

/* Voxel down-sample behind ModelFreeCollisionDetector.__init__ (utils/collision_detector.py:32-35 calls
 * open3d's PointCloud.voxel_down_sample; restated, parity unpinned).  points[n,3] f64 -> out[<=n,3] f64
 * (fp64 mean per occupied voxel, accumulated in point order, rows in ascending voxel-key order),
 * *nvox (device int) = number of rows.  workspace: b200_voxel_down_sample_workspace(n) bytes. */
size_t b200_voxel_down_sample_workspace(int n);
int b200_voxel_down_sample(const double *points, int n, double voxel_size, double *out, int32_t *nvox,
                           void *workspace, size_t workspace_bytes, void *stream);
/* The same; *nbad (device int, may be NULL) = number of input points LEFT OUT of the result because they are not
 * finite or lie outside 2^21 voxels per axis from the cloud's minimum corner (open3d raises "voxel_size is too small"
 * in that case; collision_detector.py:32-35 would propagate it).  b200_voxel_down_sample leaves them out silently. */
int b200_voxel_down_sample_checked(const double *points, int n, double voxel_size, double *out, int32_t *nvox,
                                   int32_t *nbad, void *workspace, size_t workspace_bytes, void *stream);

/* GraspGroup.nms (graspnetAPI -> grasp_nms.nms_grasp; call sites demo_pointcloud.py:249,
 * grasp_service.py:276).  grasps[M,17] f32; keep[M] i32 receives the surviving
 * row numbers in descending-score order, *nkeep (device int) their count.
 * workspace: b200_grasp_nms_workspace(m) bytes of device scratch. */
size_t b200_grasp_nms_workspace(int m);
/* The two third-party pieces of the path (graspnetAPI's grasp_nms, open3d's voxel_down_sample) are RESTATED: their
 * sources are not in the reference tree, no reference test or fixture pins them (parity unpinned).  Every choice the
 * restatement had to make is a switch; 0 = the documented default.  Process-wide, applies to the launches that follow
 * (all b200_grasp_nms* entry points / b200_voxel_down_sample*); the CPU oracle takes the same flags. */
#define B200_NMS_TRANS_INCLUSIVE 1u  /* suppress if distance <= translation_thresh        (default: <)              */
#define B200_NMS_ROT_INCLUSIVE 2u    /* ... and angle <= rotation_thresh                  (default: <)              */
#define B200_NMS_SAME_OBJECT 4u      /* only grasps with equal object id (column 16)      (default: ids ignored)    */
#define B200_NMS_FP32 8u             /* pair arithmetic in fp32                           (default: fp64)           */
#define B200_NMS_ACOS 16u            /* compare acos(clip((tr - 1) / 2)) with the angle   (default: cosines)        */
#define B200_VOXEL_ORIGIN_AT_MIN 1u  /* voxel grid origin = min_bound                     (default: - voxel / 2)    */
#define B200_VOXEL_MULTI_CTA 2u      /* not a choice of the restatement: always take the multi-launch path (library sort)
                                      * instead of the single-CTA kernel for clouds of <= 32768 points; same output */
void b200_grasp_nms_set_options(unsigned flags);
void b200_voxel_set_options(unsigned flags);

int b200_grasp_nms(const float *grasps, int m, double translation_thresh, double cos_rotation_thresh,
                   int32_t *keep, int32_t *nkeep, void *workspace, size_t workspace_bytes,
                   void *stream);

/* b200_grasp_nms without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving
 * loops, harnesses that forbid it).  The scratch is a file-scope `static __device__` array inside the library, sized
 * for m <= B200_NMS_STATIC_MAX_M grasps (532,736 bytes, held by every GPU that loads the library until exit);
 * m beyond that returns B200_ESTATIC and launches nothing.  The entry point calls no cudaMalloc / cudaFree / allocating
 * routine; inputs and outputs are the caller's, as for b200_grasp_nms, and the results are identical.  All calls on a
 * device share the one array, which keeps the previous call's contents (nothing in it needs to start at zero): calls
 * that overlap on the GPU -- two streams, concurrent graph replays -- RACE; issue them on one stream. */
#define B200_NMS_STATIC_MAX_M 2048
int b200_grasp_nms_static(const float *grasps, int m, double translation_thresh, double cos_rotation_thresh,
                          int32_t *keep, int32_t *nkeep, void *stream);

/* The whole sequence after pred_decode in one call (five graph nodes instead of nine):
 *   gg = gg[~mfcdetector.detect(gg, approach_dist, collision_thresh)]; gg = gg.nms()
 * (test.py:93-96, demo_pointcloud.py:228-249; = b200_collision_mask_device + b200_compact_rows +
 * b200_grasp_nms_device + b200_gather_rows, same arithmetic, bit-identical results).  scene: (n_max, 3) fp64 voxel
 * centres of which *n_dev are valid (n_dev may be NULL); grasps: (m_max, 17) fp32 rows of which *m_dev are valid,
 * m_max <= B200_POST_CHAIN_MAX_M (else B200_ECAPACITY: use the separate calls).  Outputs: collision_mask (m_max) u8,
 * kept_rows (m_max, 17) in score order with *nkept rows valid, and (optional) *nrows_out = the valid row count. */
#define B200_POST_CHAIN_MAX_M 2048
size_t b200_grasp_post_chain_workspace(int m_max);
int b200_grasp_post_chain(const double *scene, const int32_t *n_dev, int n_max, const float *grasps,
                          const int32_t *m_dev, int m_max, double voxel_size, double approach_dist,
                          double collision_thresh, double translation_thresh, double cos_rotation_thresh,
                          uint8_t *collision_mask, float *kept_rows, int32_t *nkept, int32_t *nrows_out,
                          void *workspace, size_t workspace_bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200GRASP_H_ */
