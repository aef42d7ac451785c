/*
 * include/geot_b200.h — C ABI of libgeot_b200.so: B200 (sm_100a) kernels for GeoT's point-cloud
 * sampling / grouping / interpolation hot path and its per-point transition-matrix path.
 *
 * This is the drop-in boundary.  The reference binds these ops through three pybind11 modules
 * (paths relative to the reference root):
 *   pointnet2_batch_cuda   openpoints/cpp/pointnet2_batch/src/pointnet2_api.cpp:10-24
 *   pointops_cuda          pointops/src/pointops_api.cpp:8-12, openpoints/cpp/pointops/src/pointops_api.cpp:13-25
 *   pointnet2._ext         pointnet2/_ext_src/src/bindings.cpp:9-21
 * Each entry point below names the reference callable it replaces.  INTEGRATION.md shows the
 * reference-side stub (ctypes / pybind) a maintainer would add.
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer on the current CUDA device; the caller owns all memory
 *     (outputs and workspaces) — nothing is allocated, freed or synchronised inside;
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*; NULL = legacy default stream);
 *   - returns GEOT_OK (0) or a negative GEOT_ERR_* code; geot_last_error() gives the message for
 *     the calling thread.  Nothing calls exit() (the reference does: sampling_gpu.cu:255-259);
 *   - tensors are contiguous fp32 / int32 in the reference's layouts; 64-bit flat indexing inside
 *     (the reference's int32 products overflow past 2^31 elements, group_points_gpu.cu:68-69);
 *   - index results are bit-exact with the reference kernels; value results match to 1e-5 relative;
 *     backward passes are deterministic (fixed summation order) where the reference uses float atomics.
 */
#ifndef GEOT_B200_H
#define GEOT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* geot_stream_t; /* cudaStream_t */

#define GEOT_OK 0
#define GEOT_ERR_INVALID (-1)     /* bad argument (NULL pointer, negative size, unsupported k ...) */
#define GEOT_ERR_CUDA (-2)        /* a CUDA launch / runtime call failed */
#define GEOT_ERR_WORKSPACE (-3)   /* workspace missing or too small */

/* Library introspection. */
int geot_abi_version(void);
const char* geot_last_error(void);
/* Number of kernels this library has launched from the calling process (bench.py's gpu_launches). */
uint64_t geot_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * Furthest point sampling.
 * flags select which of the reference's three kernels is reproduced bit-for-bit:
 *   0                         pointnet2_batch  furthest_point_sampling_wrapper   (sampling.cpp:39-48,
 *                             sampling_gpu.cu:101-216): first index 0, tie order of a 1024-thread tree
 *   GEOT_FPS_CAP512|MAGSKIP   pointnet2._ext furthest_point_sampling (pointnet2/_ext_src/src/
 *                             sampling_gpu.cu:73-176): 512-thread tree, points with |p|^2 <= 1e-3 skipped
 * `temp` mirrors the reference's caller-made (B,N) scratch pre-filled with 1e10.  The fast path keeps
 * the running min-distances on chip and never touches it (may be NULL); the large-N fallback uses it
 * as scratch and then requires it, pre-filled with 1e10 exactly like the reference.
 * --------------------------------------------------------------------------------------------- */
#define GEOT_FPS_CAP512 1
#define GEOT_FPS_MAGSKIP 2
#define GEOT_FPS_FORCE_GENERIC 4 /* testing: take the large-N fallback regardless of N */
#define GEOT_FPS_NO_PRUNE 8      /* testing / A-B timing: the un-pruned one-CTA-per-cloud kernel (every round visits every point) */
#define GEOT_FPS_SPEC 16         /* A-B: the pruned kernel with the manager guessing one round ahead (same indices) */

/* replaces pointnet2_batch_cuda.furthest_point_sampling_wrapper(B,N,M,xyz,temp,idx) and
 * pointnet2._ext.furthest_point_sampling(points, nsamples).  xyz (B,N,3) -> idx (B,M) int32. */
int geot_furthest_point_sampling(int B, int N, int M, const float* xyz, float* temp, int* idx,
                                 int flags, geot_stream_t stream);

/* replaces pointops_cuda.furthestsampling_cuda(b,n_max,xyz,offset,new_offset,tmp,idx) and, with
 * weights != NULL, furthestsampling_weights_cuda (pointops/src/sampling/sampling_cuda_kernel.cu:15-129,
 * :175-307).  xyz (n,3); offset/new_offset cumulative int32 (b) on the DEVICE; idx (new_offset[b-1])
 * receives GLOBAL indices.  n_max (largest segment) picks the tie rule exactly like the reference;
 * m_max = largest number of samples of any segment (host-known, sizes nothing but validation).
 * tmp (n) pre-filled 1e10: used only by the fallback / weighted path, may be NULL otherwise. */
int geot_furthestsampling(int b, int n_max, const float* xyz, const int* offset,
                          const int* new_offset, const float* weights, float* tmp, int* idx,
                          int flags, geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Ball query.  replaces pointnet2_batch_cuda.ball_query_wrapper(B,N,M,radius,nsample,new_xyz,xyz,idx)
 * (ball_query.cpp:29-39, ball_query_gpu.cu:15-51) and pointnet2._ext.ball_query(new_xyz,xyz,radius,nsample).
 * new_xyz (B,M,3), xyz (B,N,3) -> idx (B,M,nsample): first nsample indices k (ascending) with
 * d2 < radius*radius, remaining slots = first hit, rows with no hit all 0.  Unlike the reference the
 * kernel writes every slot itself (no caller zero-fill needed).
 * workspace: geot_ball_query_workspace_bytes(B,N) bytes of scratch (uniform-grid binning of each cloud) or
 * NULL.  With a workspace the library bins the clouds and, per cloud, takes the grid path when the ball is
 * small against the cloud; without one (or N > 65536) every cloud is scanned in index order.  Same result.
 * --------------------------------------------------------------------------------------------- */
#define GEOT_BQ_FORCE_GRID 1 /* testing: take the grid path for every cloud, whatever the cost model says */
/* The binning of the clouds depends on (xyz, radius, nsample) only, not on the query centres.  A caller that computes the
 * centres from the same cloud first (PointNet++: furthest_point_sample -> gather -> ball_query) can run it meanwhile on
 * another stream: call once with GEOT_BQ_BUILD_ONLY (new_xyz / idx may be NULL, M is ignored), order the streams, then
 * call with GEOT_BQ_PREBUILT, the SAME xyz, N, radius, nsample, workspace and grid flags.  Same indices as one plain call. */
#define GEOT_BQ_BUILD_ONLY 2
#define GEOT_BQ_PREBUILT 4
/* testing / A-B: GEOT_BQ_LEVELS(n), n in {1,2,4,8}: bin the points in (at most) n index levels instead of the count the
 * occupancy suggests (flags bits 8..11; 0 = automatic).  Same result for every n. */
#define GEOT_BQ_LEVELS(n) (((n) & 15) << 8)
#define GEOT_BQ_LEVELS_OF(flags) (((flags) >> 8) & 15)
size_t geot_ball_query_workspace_bytes(int B, int N);
int geot_ball_query(int B, int N, int M, float radius, int nsample, const float* new_xyz,
                    const float* xyz, int* idx, void* workspace, size_t workspace_bytes, int flags,
                    geot_stream_t stream);

/* No counterpart in the reference: added for callers that cannot
 * allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * geot_ball_query with a workspace, but the grid-binning scratch is a device array built into the library
 * (GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES, held by every GPU that loads the library until exit); the call
 * allocates and frees nothing.  Limit: N <= 65536 and geot_ball_query_workspace_bytes(B, N) <=
 * GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES (e.g. 16 clouds x 16384 points, or 7 x 65536); beyond it the call
 * returns GEOT_ERR_WORKSPACE with a message and launches nothing.  All calls on one GPU share the array:
 * calls that can overlap on the GPU (different streams, no ordering between them) race on it — order them.
 * The array keeps the previous call's contents; nothing in it needs clearing between calls (each call rebuilds
 * every cloud's bins before reading them). */
#define GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES (8u << 20)
int geot_ball_query_static(int B, int N, int M, float radius, int nsample, const float* new_xyz,
                           const float* xyz, int* idx, int flags, geot_stream_t stream);

/* replaces pointops_cuda.ballquery_cuda(m,radius,nsample,xyz,new_xyz,offset,new_offset,idx)
 * (openpoints/cpp/pointops/src/ballquery/ballquery_cuda_kernel.cu:26-78).  b = number of segments. */
int geot_ballquery_offset(int b, int m, float radius, int nsample, const float* xyz,
                          const float* new_xyz, const int* offset, const int* new_offset, int* idx,
                          geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * kNN query.  replaces pointops_cuda.knnquery_cuda(m,nsample,xyz,new_xyz,offset,new_offset,idx,dist2)
 * (pointops/src/knnquery/knnquery_cuda_kernel.cu:65-108).  Global int32 indices, ascending d2;
 * exact-tie cases reproduce the reference's heap order.  b = number of segments, 1 <= nsample <= 1000.
 * With offset == new_offset == NULL the inputs are batched: b clouds of n_per xyz points and m_per
 * queries each (the layout of pointops.knn / knn_cuda.KNN / knn_point), indices stay GLOBAL.
 * In the offset layout m_per must be an upper bound of the queries of any one segment (m always is).
 * n = total number of xyz points (offset[b-1] in the offset layout, b*n_per batched), m = total queries.
 * workspace: geot_knnquery_workspace_bytes(b, n, m) bytes of scratch (per-query redo flags; for segments
 * of >= 3072 points a uniform grid per segment: cell table + points sorted by cell).
 * flags: GEOT_KNN_NO_GRID forces the brute-force scan for every segment (testing / A-B timing; same results).
 * --------------------------------------------------------------------------------------------- */
#define GEOT_KNN_NO_GRID 1
/* Equal distances rank by index (lowest first) instead of by the reference kernel's heap order: the definition this
 * project uses for the un-vendored knn_cuda.KNN and for openpoints' torch knn_point (cdist + topk, whose tie order is
 * unspecified).  Skips the replay of the queries that have an exact tie among their k+1 best (one such query of 32000
 * costs 0.2 ms: it walks its whole segment through a sequential heap).  Needs nsample <= 127. */
#define GEOT_KNN_TIES_BY_INDEX 2
size_t geot_knnquery_workspace_bytes(int b, int n, int m);
int geot_knnquery(int b, int n, int m, int nsample, const float* xyz, const float* new_xyz,
                  const int* offset, const int* new_offset, int n_per, int m_per, int* idx,
                  float* dist2, void* workspace, size_t workspace_bytes, int flags, geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Grouping / gather (channel-major, batched).
 * --------------------------------------------------------------------------------------------- */
/* replaces pointnet2_batch_cuda.group_points_wrapper(B,C,N,M,ns,points,idx,out) (group_points_gpu.cu:53-72)
 * and pointnet2._ext.group_points.  points (B,C,N), idx (B,M,ns) -> out (B,C,M,ns). */
int geot_group_points(int B, int C, int N, int M, int nsample, const float* points, const int* idx,
                      float* out, geot_stream_t stream);
/* Point-major gather: src (B,N,D), idx (B,M) -> out (B,M,D) = src[b, idx[b,j], :].  D = 3 gives the sampled centres
 * new_xyz = xyz[fps_idx] that the reference computes as gather_operation(xyz.transpose(1,2).contiguous(), idx)
 * .transpose(1,2).contiguous() (sampling_gpu.cu:15-31 between two transpose copies). */
int geot_gather_rows(int B, int N, int M, int D, const float* src, const int* idx, float* out, geot_stream_t stream);
/* Fused xyz grouping of QueryAndGroup / KNNGroup (group.py:235-250, :297-312; pointnet2_utils.py:291-373):
 * out[b,c,j,s] = (xyz[b,idx[b,j,s],c] - new_xyz[b,j,c]) / divisor, straight from the (B,N,3) cloud - replaces
 * transpose(1,2).contiguous() + grouping_operation + subtraction (+ division).  relative = 0 skips the subtraction,
 * divisor = 0 skips the division.  xyz (B,N,3), new_xyz (B,M,3), idx (B,M,ns) -> out (B,3,M,ns); bit-identical to
 * the composite. */
int geot_group_xyz_relative(int B, int N, int M, int nsample, const float* xyz, const float* new_xyz,
                            const int* idx, int relative, float divisor, float* out, geot_stream_t stream);
/* testing / A-B timing (a THREAD-LOCAL override, like the error slot: it affects only later calls of the calling thread;
 * results are identical): mode 1 = always the direct kernels (L2 gathers,
 * brute-force three_nn, warp-per-point transition_apply, ...), mode 2 = the staged / grid kernels wherever their hard
 * limits allow, ignoring the work-size heuristics, mode 3 = three_interpolate forward without its channel-interleaved
 * kernel (the plain row-staged one instead), mode 0 = default dispatch.  Modes 4-9 switch single kernel variants off or
 * on for A/B runs (4 three_interpolate without the persistent kernel, 5 / 8 its backward never / always streamed,
 * 6 / 7 manifold loss on the warp / tile kernels, 9 grouping backward waiting for whole gradient rows). */
void geot_group_points_config(int mode);
/* replaces pointnet2_batch_cuda.gather_points_wrapper(B,C,N,M,points,idx,out) (sampling_gpu.cu:15-31)
 * and pointnet2._ext.gather_points.  points (B,C,N), idx (B,M) -> out (B,C,M). */
int geot_gather_points(int B, int C, int N, int M, const float* points, const int* idx, float* out,
                       geot_stream_t stream);

/* Deterministic scatter-add plan shared by every backward pass that the reference does with float
 * atomics: a CSR of source positions grouped by target, in ascending source order.
 *   idx (B,E) int32 with values in [0,N)  ->  row_ptr (B,N+1) int32, perm (B,E) int32.
 * Workspace layout: [row_ptr | perm]; query the size first.  Reusable across channels and calls that
 * share idx (e.g. the xyz and feature grouping of one QueryAndGroup). */
size_t geot_scatter_plan_bytes(int B, int N, int64_t E);
int geot_scatter_plan_build(int B, int N, int64_t E, const int* idx, void* plan, size_t plan_bytes,
                            geot_stream_t stream);

/* replaces pointnet2_batch_cuda.group_points_grad_wrapper(B,C,N,M,ns,grad_out,idx,grad_points)
 * (group_points_gpu.cu:14-31, atomicAdd :30) and pointnet2._ext.group_points_grad.
 * grad_out (B,C,M,ns) -> grad_points (B,C,N), every element written (no caller zero-fill needed).
 * plan = geot_scatter_plan_build(B,N,M*ns,idx). */
int geot_group_points_grad(int B, int C, int N, int M, int nsample, const float* grad_out,
                           const void* plan, float* grad_points, geot_stream_t stream);
/* Large neighbourhood tensors (M*nsample > 65535, or a row that does not fit shared memory): the gradient row is
 * processed in `pieces` contiguous parts through the shared-memory kernel.  pieces = geot_scatter_plan_pieces(N, E)
 * (1 = use the call above); the plan must then be built as geot_scatter_plan_build(B*pieces, N, E/pieces, idx, ...),
 * i.e. one CSR per (cloud, piece) over the SAME idx buffer.  Result = sum over pieces in order: deterministic. */
int geot_scatter_plan_pieces(int N, int64_t E);
int geot_group_points_grad_pieces(int B, int C, int N, int M, int nsample, int pieces, const float* grad_out,
                                  const void* plan, float* grad_points, geot_stream_t stream);
/* replaces pointnet2_batch_cuda.gather_points_grad_wrapper (sampling_gpu.cu:53-70). plan over E = M. */
int geot_gather_points_grad(int B, int C, int N, int M, const float* grad_out, const void* plan,
                            float* grad_points, geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * three_nn / three_interpolate (batched).
 * --------------------------------------------------------------------------------------------- */
/* replaces pointnet2_batch_cuda.three_nn_wrapper(B,n,m,unknown,known,dist2,idx) (interpolate_gpu.cu:16-59)
 * and pointnet2._ext.three_nn.  unknown (B,n,3), known (B,m,3) -> dist2 (B,n,3) SQUARED, idx (B,n,3). */
int geot_three_nn(int B, int n, int m, const float* unknown, const float* known, float* dist2,
                  int* idx, geot_stream_t stream);
/* replaces three_interpolate_wrapper(B,C,m,n,points,idx,weight,out) (interpolate_gpu.cu:84-104). */
int geot_three_interpolate(int B, int C, int m, int n, const float* points, const int* idx,
                           const float* weight, float* out, geot_stream_t stream);
/* replaces three_interpolate_grad_wrapper(B,C,n,m,grad_out,idx,weight,grad_points)
 * (interpolate_gpu.cu:127-149, 3 atomicAdds).  plan = geot_scatter_plan_build(B, m, 3*n, idx). */
int geot_three_interpolate_grad(int B, int C, int n, int m, const float* grad_out, const float* weight,
                                const void* plan, float* grad_points, geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Offset-format (channel-last) ops of pointops_cuda (openpoints/cpp/pointops/src).
 * --------------------------------------------------------------------------------------------- */
/* grouping_forward_cuda(m,nsample,c,input,idx,output)  grouping/grouping_cuda_kernel.cu:5-14 */
int geot_grouping_forward(int m, int nsample, int c, const float* input, const int* idx,
                          float* output, geot_stream_t stream);
/* grouping_backward_cuda (:16-25, atomicAdd).  n = rows of grad_input; plan over (B=1, N=n, E=m*nsample). */
int geot_grouping_backward(int m, int nsample, int c, int n, const float* grad_output,
                           const void* plan, float* grad_input, geot_stream_t stream);
/* interpolation_forward_cuda(n,c,k,input,idx,weight,output)  interpolation/interpolation_cuda_kernel.cu:5-18.
 * Writes output (the reference accumulates into a caller-zeroed buffer; same values). */
int geot_interpolation_forward(int n, int c, int k, const float* input, const int* idx,
                               const float* weight, float* output, geot_stream_t stream);
/* interpolation_backward_cuda (:20-33).  m = rows of grad_input; plan over (1, m, n*k). */
int geot_interpolation_backward(int n, int c, int k, int m, const float* grad_output,
                                const float* weight, const void* plan, float* grad_input,
                                geot_stream_t stream);
/* subtraction_forward_cuda / subtraction_backward_cuda  subtraction/subtraction_cuda_kernel.cu:5-30.
 * plan over (1, n, n*nsample). */
int geot_subtraction_forward(int n, int nsample, int c, const float* input1, const float* input2,
                             const int* idx, float* output, geot_stream_t stream);
int geot_subtraction_backward(int n, int nsample, int c, const float* grad_output, const void* plan,
                              float* grad_input1, float* grad_input2, geot_stream_t stream);
/* aggregation_forward_cuda / aggregation_backward_cuda  aggregation/aggregation_cuda_kernel.cu:5-39.
 * plan over (1, n, n*nsample). */
int geot_aggregation_forward(int n, int nsample, int c, int w_c, const float* input,
                             const float* position, const float* weight, const int* idx,
                             float* output, geot_stream_t stream);
int geot_aggregation_backward(int n, int nsample, int c, int w_c, const float* input,
                              const float* position, const float* weight, const int* idx,
                              const float* grad_output, const void* plan, float* grad_input,
                              float* grad_position, float* grad_weight, geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Instance-dependent transition-matrix path (GeoT's noisy-pseudo-label loss).
 * All tensors fp32 unless noted; C = number of classes (2 <= C <= 64).
 * --------------------------------------------------------------------------------------------- */
/* replaces sig_t_mean.forward (openpoints/models/backbone/transformer.py:1120-1131, a Python loop of C
 * Linear(2C->C) calls + clamp + F.normalize) and, with cm == NULL, sig_t.forward (:1089-1097).
 *   x (B,C,N) class probabilities; W: sig_t_mean (C, C, 2C) = stacked fc[kk].weight, w_row_stride = 2C;
 *   sig_t (C*C, C) = fc.weight, w_row_stride = C; cm (C,C) or NULL  ->  insT (B*N, C, C).
 * workspace: geot_sigt_workspace_bytes(C) (the point-independent bias W[:, :, C:] . cm). */
size_t geot_sigt_workspace_bytes(int C);
int geot_sigt_forward(int B, int C, int N, const float* x, const float* W, int w_row_stride,
                      const float* cm, float* insT, void* workspace, size_t workspace_bytes,
                      geot_stream_t stream);
/* backward of the above w.r.t. W (autograd of transformer.py:1123-1129): grad_insT (B*N,C,C) ->
 * grad_W (same shape as W; every element written).  g_raw (B*N, C*C) receives the gradient w.r.t. the
 * pre-clamp products (the caller may derive grad_x = g_raw @ W[:, :, :C] from it; the reference call site
 * detaches x, train.py:548).  Deterministic: fixed-order two-stage reduction over points. */
size_t geot_sigt_backward_workspace_bytes(int B, int C, int N);
int geot_sigt_backward(int B, int C, int N, const float* x, const float* W, int w_row_stride,
                       const float* cm, const float* grad_insT, float* g_raw, float* grad_W,
                       void* workspace, size_t workspace_bytes, geot_stream_t stream);

/* replaces train.py:549-552: newT = L1norm_rows(lam*ema_t_corr + (1-lam)*insT); out[b,:,n] = logits[b,:,n] . newT[bN+n]
 * (blend + F.normalize + bmm + two permutes in one pass over insT).  logits/out (B,C,N). */
int geot_transition_apply_forward(int B, int C, int N, const float* logits, const float* ema_t,
                                  const float* insT, float lam, float* out, geot_stream_t stream);
/* its backward: grad_out (B,C,N) -> grad_logits (B,C,N) and grad_insT (B*N,C,C); either may be NULL. */
int geot_transition_apply_backward(int B, int C, int N, const float* logits, const float* ema_t,
                                   const float* insT, float lam, const float* grad_out,
                                   float* grad_logits, float* grad_insT, geot_stream_t stream);

/* replaces threeD_space_loss.forward / feature_space_loss.forward after their kNN (utils/insT_loss.py:72-110,
 * :22-58): k rounds of index_select + cat growing (BN,k,C,C) tensors become one gather pass.
 *   points (B*N, D) positions (D=3) or logits (D=C); labels (B*N) int64; insT (B*N, C*C);
 *   idx (B,N,k) int32 LOCAL neighbour indices (self excluded); mode GEOT_MANIFOLD_3D: w in {0,1}*exp, per point
 *   sum(w*Td)/(sum(w)+1e-3), mean over points; GEOT_MANIFOLD_FEATURE: w in {-1,+1}*exp, mean over points*k.
 * Outputs: coef (B*N,k) = d loss_point / d Tdist (kept for the backward), loss (1 float, device).
 * workspace: geot_manifold_workspace_bytes(B*N).
 * order (B,N) int32 or NULL: the order in which the points of each cloud are PROCESSED (geot_spatial_order: Morton
 * cells of the first three coordinates).  It keeps a neighbourhood's matrices L2-resident - the (BN,C,C) tensor does
 * not fit the L2 - and changes no result bit: every point writes its own outputs, the final sum runs in index order. */
#define GEOT_MANIFOLD_3D 0
#define GEOT_MANIFOLD_FEATURE 1
size_t geot_manifold_workspace_bytes(int64_t BN);
int geot_spatial_order(int B, int N, int D, const float* points, int* order, geot_stream_t stream);
int geot_manifold_loss_forward(int B, int N, int D, int C, int k, const float* points,
                               const int64_t* labels, const float* insT, const int* idx, const int* order,
                               float sigma, int mode, float* coef, float* loss, void* workspace,
                               size_t workspace_bytes, geot_stream_t stream);
/* backward w.r.t. insT (weights are detached in the reference, insT_loss.py:106-108): grad_insT (B*N,C*C),
 * every element written, fixed summation order.  plan = geot_scatter_plan_build(B, N, N*k, idx);
 * grad_loss: 1 float on the device (upstream gradient of the scalar loss). */
int geot_manifold_loss_backward(int B, int N, int C, int k, const float* insT, const int* idx,
                                const float* coef, const void* plan, const int* order, const float* grad_loss,
                                float* grad_insT, geot_stream_t stream);

/* replaces the per-class loop train.py:509-526 (filter_outlier off): for every class cc the FIRST point in
 * (b,n) order with the largest pred_u[b,cc,n]; class_T[cc] = pred_u[b,:,n].  best_val (C), best_idx (C) int64
 * flat b*N+n are returned so ranks can combine their picks (arg-max over ranks, then a row broadcast).
 * workspace: geot_class_pick_workspace_bytes(C). */
size_t geot_class_pick_workspace_bytes(int C);
int geot_class_pick(int B, int C, int N, const float* pred_u, float* class_T, float* best_val,
                    int64_t* best_idx, void* workspace, size_t workspace_bytes, geot_stream_t stream);
/* replaces train.py:507,528-545,556-557 (Gaussian geometric prior over label_proj, blend, the reference's
 * keepdim-less normalisation copied verbatim, EMA): class_T, ema_t (C,C), sigma (C), label_proj (C) int32
 * -> ema_t_corr (C,C), ema_t_out (C,C; may alias ema_t). */
int geot_class_t_step(int C, const float* class_T, const float* ema_t, const float* sigma,
                      const int* label_proj, float geo_lambda, float ema_decay, float* ema_t_corr,
                      float* ema_t_out, geot_stream_t stream);
/* replaces the inner loop of cal_mean_feature (train.py:884-895), bug-for-bug (probs[target] indexes rows by
 * class id): probs (BN,C), target (BN) int64; cm (C,C) and c_num (C) updated in place.
 * workspace: geot_class_pick_workspace_bytes(C). */
int geot_mean_feature_update(int64_t BN, int C, const float* probs, const int64_t* target, float* cm,
                             float* c_num, void* workspace, size_t workspace_bytes,
                             geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Unsupervised poly-1 focal loss on the transition-corrected logits.  replaces Poly1FocalLoss_U_corr.forward
 * (openpoints/loss/build.py:830-892; criterion_u of cfgs/tooth_semi/transformer_finetune_fixmatch_ntm.yaml:54,
 * called at examples/segmentation/train.py:592) and its autograd, weight = pos_weight = None.
 * logits (B,C,N), labels (B,N) int64, conf (B,N) teacher confidence compared >= thresh, or mask (B,N) float weights
 * (non-NULL mask wins) -> loss[0] = sum(keep * poly1) / (C * sum(keep) + 0.001), denom[0] = that denominator.
 * backward: grad_logits (B,C,N) = grad_loss[0] * keep * d poly1 / d logit / denom, every element written.
 * workspace: geot_poly1_focal_workspace_bytes(B*N) (per-CTA partial sums; summed in a fixed order).
 * --------------------------------------------------------------------------------------------- */
size_t geot_poly1_focal_workspace_bytes(int64_t BN);
int geot_poly1_focal_forward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                             const float* mask, float thresh, float epsilon, float alpha, float gamma,
                             float* loss, float* denom, void* workspace, size_t workspace_bytes, geot_stream_t stream);
int geot_poly1_focal_backward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                              const float* mask, float thresh, float epsilon, float alpha, float gamma,
                              const float* denom, const float* grad_loss, float* grad_logits, geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * DGCNN edge features.  replaces the tensor algebra of DGCNN_Propagation.get_graph_feature after
 * its kNN (openpoints/models/backbone/transformer.py:355-363: transpose, row gather, permute, expand, subtract, cat).
 * x_k (B,C,Nk) key features, x_q (B,C,Nq) query features, idx (B,Nq,k) int32 neighbour indices into the Nk axis (cloud-
 * local; the layout the kNN kernel writes) -> out (B,2C,Nq,k): out[:, :C] = x_k[.., idx] - x_q, out[:, C:] = x_q.
 * geot_graph_feature_grad_q: grad_out (B,2C,Nq,k) -> grad_x_q (B,C,Nq) = sum_j (grad_out[:, C:] - grad_out[:, :C]);
 * grad_x_k is geot_group_points_grad of grad_out[:, :C] over the same idx (deterministic scatter plan).
 * --------------------------------------------------------------------------------------------- */
int geot_graph_feature(int B, int C, int Nk, int Nq, int k, const float* x_k, const float* x_q, const int* idx,
                       float* out, geot_stream_t stream);
int geot_graph_feature_grad_q(int B, int C, int Nq, int k, const float* grad_out, float* grad_x_q,
                              geot_stream_t stream);

/* Poly1FocalLoss_U_T.forward (openpoints/loss/build.py:595-688, selected at examples/segmentation/train.py:594-596) and its
 * autograd: the same poly-1 focal terms on the UNcorrected student logits, every point scaled by
 * beta = conf / pred_t[b, label, n] (pred_t (B,C,N) = the transition-corrected prediction, build.py:661-670; conf =
 * logits_pred (B,N), required even when mask is given - the mask then only replaces the keep weights).
 * backward writes every element of grad_logits and grad_pred_t (B,C,N; non-zero only in the label row). */
int geot_poly1_focal_t_forward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                               const float* mask, const float* pred_t, float thresh, float epsilon, float alpha,
                               float gamma, float* loss, float* denom, void* workspace, size_t workspace_bytes,
                               geot_stream_t stream);
int geot_poly1_focal_t_backward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                                const float* mask, const float* pred_t, float thresh, float epsilon, float alpha,
                                float gamma, const float* denom, const float* grad_loss, float* grad_logits,
                                float* grad_pred_t, geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Validation upsampling + metrics.
 * geot_interp_weights: replaces `dist_recip = 1.0 / (dist + 1e-8); norm = sum(dist_recip, 2); weight = dist_recip / norm`
 * (examples/segmentation/train.py:794-796, openpoints/models/layers/upsampling.py:98-100,
 * pointnet2/pointnet2_modules.py:619-623) fed with the SQUARED distances geot_three_nn returns: total = B*n targets,
 * dist2 (total,3) -> weight (total,3) and, when dist != NULL, dist = sqrt(dist2) (what the reference's three_nn returns).
 * geot_interp_argmax: replaces `three_interpolate(logit, idx, weight).argmax(dim=1)` (train.py:797-798): points (B,C,m),
 * idx/weight (B,n,3) -> pred (B,n) int64 = first arg-max over the C blended rows; score (B,n) = that maximum, may be NULL.
 * geot_seg_counts: replaces the per-class logical_and/logical_or sums and the accuracy count of get_seg_metrics
 * (train.py:817-826): pred, label (n) int64 -> counts (3C+2) int64 = inter[C] | pred_cnt[C] | label_cnt[C] | correct |
 * vertices with pred or label outside [0,C); union = pred_cnt + label_cnt - inter.  counts is zeroed by the call.
 * --------------------------------------------------------------------------------------------- */
int geot_interp_weights(int64_t total, const float* dist2, float* dist, float* weight, geot_stream_t stream);
int geot_interp_argmax(int B, int C, int m, int n, const float* points, const int* idx, const float* weight,
                       int64_t* pred, float* score, geot_stream_t stream);
int geot_seg_counts(int64_t n, int C, const int64_t* pred, const int64_t* label, int64_t* counts,
                    geot_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Multi-GPU statistics exchange (no reference counterpart: the reference is single-process,
 * examples/segmentation/train.py:7, and reduces nothing).  One packed record per rank around ONE all-gather:
 *   record = [ best_val (C) | class_T (C*C) | cm * c_num (C*C) | c_num (C) | grads (P) ],  geot_stats_record_floats(C,P)
 * geot_stats_pack writes this rank's record; geot_stats_combine reads the W gathered records (W, L) and writes the
 * global class-anchor rows (arg-max of best_val over ranks, lowest rank on ties = the single-process first-max of
 * train.py:509-526 over the concatenated batch), the count-weighted class means / summed counts of
 * cal_mean_feature (train.py:884-895) and grads = grad_scale * sum over ranks, all in fixed rank order (every rank
 * computes identical bits).  grads may be NULL in both calls (P floats are then zero-filled / ignored).
 * --------------------------------------------------------------------------------------------- */
int64_t geot_stats_record_floats(int C, int64_t P);
int geot_stats_pack(int C, int64_t P, const float* class_T, const float* best_val, const float* cm, const float* c_num,
                    const float* grads, float* record, geot_stream_t stream);
int geot_stats_combine(int W, int C, int64_t P, const float* gathered, float grad_scale, float* class_T, float* cm,
                       float* c_num, float* grads, geot_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GEOT_B200_H */
