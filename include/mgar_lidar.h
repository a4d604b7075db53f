/*
 * mgar_lidar.h - C ABI of libmgar_lidar.so: the B200 (sm_100a) LiDAR hot path of MGAR-net.
 *
 * Every entry point replaces one function the reference binds through pybind11 in its
 * pcdet/ops extensions (citations: file:line under /root/reference/pcdet/ops/).  Plain
 * pointers and sizes only; no torch types.  Unless a name ends in _host, every pointer is
 * a DEVICE pointer, outputs are caller-allocated (the reference's convention),
 * and work is enqueued on `stream` (a cudaStream_t passed as void*; NULL = legacy default).
 *
 * Return value: 0 on success, negative MGAR_E* on failure; mgar_last_error() returns a
 * thread-local message.  (The reference prints and calls exit(-1); we never do.)
 * There is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef MGAR_LIDAR_H
#define MGAR_LIDAR_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGAR_OK 0
#define MGAR_EINVAL (-1)  /* bad argument (shape, null pointer, unsupported size) */
#define MGAR_ECUDA (-2)   /* CUDA runtime / launch error                           */
#define MGAR_ENOMEM (-3)  /* workspace too small / allocation failure              */

#define MGAR_ABI_VERSION 2

#if defined(__GNUC__)
#define MGAR_API __attribute__((visibility("default")))
#else
#define MGAR_API
#endif

MGAR_API int mgar_abi_version(void);
MGAR_API const char *mgar_last_error(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
MGAR_API unsigned long long mgar_launch_count(void);

/* Scratch of the stack-layout (ragged) operators.  The reference's kernels re-derive a point's batch with an O(B) loop
 * per thread (e.g. ball_query_gpu.cu:27-35); here one small launch turns the (B) count arrays into exclusive prefix
 * arrays that the kernels binary-search.  Those arrays live in CALLER-PROVIDED device scratch of
 * mgar_stack_workspace_bytes(B) bytes, passed as (workspace, workspace_bytes) to every entry that takes batch counts
 * (NULL / 0 with the uniform layout).  mgar_sa_mlp_fwd_train / _bwd take it BEHIND their own workspace: their
 * workspace must then hold round_up(mgar_sa_mlp_train_workspace_bytes_for(..), 16) + mgar_stack_workspace_bytes(B).
 * The library never allocates device memory. */
MGAR_API size_t mgar_stack_workspace_bytes(int B);

/* Pinned host staging for the host-buffer front end: cudaHostAlloc (portable; write_combined != 0 adds
 * cudaHostAllocWriteCombined - for buffers the CPU only writes and the copy engines read) / cudaFreeHost. */
MGAR_API int mgar_host_alloc(void **ptr, size_t bytes, int write_combined);
MGAR_API int mgar_host_free(void *ptr);

/* ---- roiaware_pool3d: in-box tests -------------------------------------------------------- */

/* points_in_boxes_gpu: roiaware_pool3d/src/roiaware_pool3d.cpp:98-118 (kernel .cu:313-336).
 * boxes (B,M,7) [cx,cy,cz,dx,dy,dz,heading], pts (B,N,3) -> box_idx (B,N) int32: index of the
 * FIRST box containing the point (margin 1e-5), else -1.  Every element is written. */
MGAR_API int mgar_points_in_boxes(const float *boxes, const float *pts, int *box_idx, int B, int M, int N,
                         void *stream);

/* (M,N) 0/1 in-box matrix on the device.  cpu_semantics=0: the GPU kernels' arithmetic
 * (margin 1e-5, FMA-contracted rotation, device cosf/sinf; roipoint_pool3d_kernel.cu:38-60,
 * transposed).  cpu_semantics=1: the arithmetic of points_in_boxes_cpu (margin 1e-2, unfused
 * rotation) and `cs` (M,2) must hold the host-computed cosf(-heading), sinf(-heading). */
MGAR_API int mgar_in_box_mask(const float *boxes, const float *pts, const float *cs, int *mask, int M, int N,
                     int cpu_semantics, void *stream);

/* points_in_boxes_cpu: roiaware_pool3d.cpp:143-168.  HOST buffers in and out: boxes (M,7),
 * pts (N,3) -> mask (M,N) int32 0/1 with the CPU function's margin (1e-2).  Runs on the GPU
 * (H2D, kernel, D2H, synchronous); only the M box headings are pre-processed on the host so the
 * result is bit-identical to the reference's glibc trig.  workspace: DEVICE staging (on the current device) of
 * mgar_points_in_boxes_mask_host_workspace_bytes(M, N) bytes, owned by the caller. */
MGAR_API size_t mgar_points_in_boxes_mask_host_workspace_bytes(int M, int N);
MGAR_API int mgar_points_in_boxes_mask_host(const float *boxes_host, const float *pts_host, int *mask_host,
                                   int M, int N, void *workspace, size_t workspace_bytes);

/* ---- roipoint_pool3d ------------------------------------------------------------------------ */

/* roipool3d_gpu: roipoint_pool3d/src/roipoint_pool3d.cpp:23-54 (kernels .cu:38-134).
 * xyz (B,N,3), boxes (B,M,7) ALREADY enlarged by the caller, feat (B,N,C) ->
 * pooled (B,M,S,3+C), empty (B,M) int32 (every element of both is written; empty boxes give
 * zeros and flag 1).  pts_idx (B,M,S) optional (may be NULL): the pooled point indices.
 * canonical != 0 additionally applies pointrcnn_head.py:121-129 (subtract centre, rotate by
 * -heading) to the xyz columns using `boxes` as the RoIs. */
MGAR_API int mgar_roipoint_pool(const float *xyz, const float *boxes, const float *feat, float *pooled,
                       int *empty, int *pts_idx, int B, int N, int M, int S, int C, int canonical,
                       void *stream);

/* Two-pass form of the same pooling for whole batches (the throughput path): pass 1 streams the points once for
 * ALL boxes of a frame and leaves an in-box bit mask, mask (B, M, mgar_box_mask_words(N)) uint32, bit (n & 31) of
 * word n >> 5 = point n inside box m of boxes_pool (margin 1e-5, the test of roipoint_pool3d_kernel.cu:38-60); the
 * buffer is zero-filled by the call.  When box_idx != NULL the same pass also writes points_in_boxes_gpu's
 * first-hit index (roiaware_pool3d_kernel.cu:313-336) for boxes_first (pass the same pointer as boxes_pool when
 * the pooling boxes are not enlarged: the tests are then shared).  Pass 2 turns the masks into the pooled rows;
 * arguments and outputs as mgar_roipoint_pool, plus pooled_cnt (B,M) int32, optional: min(in-box points, S) - the
 * number of DISTINCT rows of each pooled cloud (the scheduling weight of mgar_fps_ball_query). */
MGAR_API size_t mgar_box_mask_words(int N);
/* workspace (optional, mgar_box_masks_workspace_bytes(B, M) bytes of device scratch, 0 when M > 255): the per-frame box
 * grid is then built once per frame by a small kernel instead of once per CTA in shared memory (same outputs). */
MGAR_API size_t mgar_box_masks_workspace_bytes(int B, int M);
MGAR_API int mgar_box_masks(const float *boxes_first, const float *boxes_pool, const float *pts, int *box_idx,
                            unsigned *mask, int B, int M, int N, void *workspace, size_t workspace_bytes, void *stream);
MGAR_API int mgar_roipoint_pool_from_masks(const float *xyz, const float *boxes, const float *feat, const unsigned *mask,
                                           float *pooled, int *empty, int *pts_idx, int *pooled_cnt, int B, int N, int M,
                                           int S, int C, int canonical, void *stream);

/* ---- roiaware_pool3d: voxel pooling inside boxes ---------------------------------------------- */

/* roiaware_pool3d_gpu / roiaware_pool3d_gpu_backward: roiaware_pool3d/src/roiaware_pool3d.cpp:20-96 (kernels
 * roiaware_pool3d_kernel.cu:36-312).  rois (boxes_num,7), pts (pts_num,3), pts_feature (pts_num,channels) ->
 * pts_idx_of_voxels (boxes_num,out_x,out_y,out_z,max_pts_each_voxel) int32 (slot 0 = count, then point indices in
 * index order; ZERO-FILLED BY THE CALLER as in roiaware_pool3d_utils.py:71-72), argmax (…,channels) int32 (max
 * pooling only; -1 = empty voxel), pooled_features (…,channels) written only for non-empty voxels (caller
 * zero-fills).  pool_method 0 = max, 1 = avg.  workspace: mgar_roiaware_pool3d_workspace_bytes(boxes_num,pts_num)
 * bytes of device scratch (the in-box bit masks; the reference cudaMallocs 32x as much per call).  out sizes are
 * limited to 256 per axis: the reference packs the voxel index in 8 bits per axis and silently corrupts larger
 * grids.  backward: grad_in (pts_num,channels) accumulated with atomics into the caller's zero-filled buffer. */
MGAR_API size_t mgar_roiaware_pool3d_workspace_bytes(int boxes_num, int pts_num);
MGAR_API int mgar_roiaware_pool3d_fwd(const float *rois, const float *pts, const float *pts_feature, int *argmax,
                                      int *pts_idx_of_voxels, float *pooled_features, void *workspace, int boxes_num,
                                      int pts_num, int channels, int max_pts_each_voxel, int out_x, int out_y,
                                      int out_z, int pool_method, void *stream);
MGAR_API int mgar_roiaware_pool3d_bwd(const int *pts_idx_of_voxels, const int *argmax, const float *grad_out,
                                      float *grad_in, int boxes_num, int out_x, int out_y, int out_z, int channels,
                                      int max_pts_each_voxel, int pool_method, void *stream);

/* ---- pointnet2_stack: farthest point sampling ------------------------------------------------- */

/* farthest_point_sampling_wrapper: pointnet2/pointnet2_stack/src/sampling.cpp:21-34
 * (kernel sampling_gpu.cu:25-140).  xyz (b,n,3), temp (b,n) pre-filled by the caller (1e10;
 * clobbered with the final min-distances) -> idx (b,m) int32.  Tie order follows the
 * reference's block reduction for block size opt_n_threads(n).
 * Extensions (the pybind mirror passes 3 / temp / NULL): xyz_stride = floats between rows (4 reads the
 * xyz columns of a pooled (b,n,4) array in place); temp may be NULL for n <= 8192 (distances start at
 * 1e10 and are not written back); new_xyz (b,m,3) optional: the sampled coordinates, gathered in-kernel. */
MGAR_API int mgar_fps_batch(int b, int n, int m, const float *xyz, int xyz_stride, float *temp, int *idx,
                            float *new_xyz, void *stream);

/* stack_farthest_point_sampling_wrapper: sampling.cpp:37-56 (kernel sampling_gpu.cu:188-319).
 * xyz (sum n_i,3), temp (sum n_i), xyz_batch_cnt (batch) int32 device, num_sampled (batch)
 * int32 device -> idx (sum m_i) GLOBAL row indices.  max_n = upper bound of n_i (host knows
 * the total N; pass N when unknown).  Block size is fixed at 1024 as in the reference. */
MGAR_API int mgar_fps_stack(int batch, int max_n, const float *xyz, float *temp, const int *xyz_batch_cnt,
                   int *idx, const int *num_sampled, void *stream);

/* ---- pointnet2_stack: ball query / grouping ----------------------------------------------------- */

/* ball_query_wrapper_stack: ball_query.cpp:29-45 (kernel ball_query_gpu.cu:16-66).
 * idx (M,nsample): batch-local indices of the first nsample points with d2 < radius^2 in index order,
 * padded with the first hit; rows without neighbours get idx[0] = -1.  fill = 0 is the reference
 * contract (idx zero-filled by the caller, the other slots of such rows are left alone); fill = 1
 * writes every slot (-1, 0, 0, ...) so idx may be uninitialised; fill = 2 leaves such rows untouched (the
 * pointnet2_batch kernel's behaviour, pointnet2_batch/src/ball_query_gpu.cu:15-51: no -1 marker).
 * Extensions: xyz_stride (floats between rows); uniform_n > 0: every cloud has uniform_n points and
 * uniform_m centroids (B * uniform_m == M) and the count arrays may be NULL. */
MGAR_API int mgar_ball_query_stack(int B, int M, float radius, int nsample, const float *new_xyz,
                          const int *new_xyz_batch_cnt, const float *xyz, int xyz_stride,
                          const int *xyz_batch_cnt, int uniform_n, int uniform_m, int *idx, int fill,
                          void *workspace, size_t workspace_bytes, void *stream);

/* voxel_query_wrapper_stack: voxel_query.cpp:25-41 (kernel voxel_query_gpu.cu:10-89).  new_xyz (M,3), xyz (N,3),
 * new_coords (M,4) int32 [batch, z, y, x] (16-byte aligned), point_indices (B,R1,R2,R3) int32 voxel -> point
 * index or -1.  idx (M,nsample): GLOBAL indices of the first nsample points with !(d2 > radius^2) met while
 * walking the (2*z_range+1)(2*y_range+1)(2*x_range+1) neighbouring voxels in (dz,dy,dx) order, padded with the
 * first hit; a query without neighbours gets idx[0] = -1 and its other slots are left alone (caller zero-fills). */
MGAR_API int mgar_voxel_query_stack(int M, int R1, int R2, int R3, int nsample, float radius, int z_range, int y_range,
                                    int x_range, const float *new_xyz, const float *xyz, const int *new_coords,
                                    const int *point_indices, int *idx, void *stream);

/* farthest_point_sampling_wrapper (sampling.cpp:21-34) followed by ball_query_wrapper_stack (ball_query.cpp:29-45)
 * around the sampled points, in ONE pass for b equally sized clouds: round j of the sampling already measures every
 * point against sample j-1, which is the distance the ball query around that centroid tests (bit-identical), so the
 * query costs one ballot per 32 points instead of a second kernel.  xyz (b,n,xyz_stride) -> fps_idx (b,m) int32,
 * new_xyz (b,m,3), ball_idx (b*m,nsample) cloud-local indices exactly as mgar_fps_batch + mgar_ball_query_stack
 * (uniform layout, fill = 1) produce them.  Supported (mgar_fps_ball_query_supported != 0): nsample <= 64 and
 * n in {128,256,512} (a CTA of four warps per cloud, any number of clouds) or n in {32,64} with b >= 592 clouds (one warp
 * per cloud); otherwise call the two entries.
 */
MGAR_API int mgar_fps_ball_query_supported(int b, int n, int m, int nsample);
/*
 * live_cnt (b) int32, optional: the number of DISTINCT samples of every cloud.  Pooled clouds repeat their points
 * (roipoint_pool3d_kernel.cu:86-97), and once every distinct point is taken the reference's arg-max returns index 0
 * (all min-distances are 0), so fps_idx[live_cnt ..] = 0 and those centroids repeat centroid 0.
 * write_dup_rows: 1 = every ball_idx row is written (the reference's output); 0 = only the rows of the centroids
 * [0, round_up(live_cnt, 4)) - for a consumer that walks the distinct centroids only (mgar_sa_mlp_fwd_tc with
 * group_live_cnt); the other rows are left untouched.
 * cloud_weight (b) int32, optional: an estimate of each cloud's sampling rounds (its number of distinct points, e.g.
 * pooled_cnt of mgar_roipoint_pool_from_masks).  The clouds are then scheduled heaviest first (a one-CTA counting sort
 * into `workspace`, mgar_fps_ball_query_workspace_bytes(b) bytes of device scratch) so that the long clouds do not
 * form a tail; results do not depend on it.  NULL: clouds run in index order, no workspace needed.
 * With n = 512 and b >= 1184 weighted clouds the call issues TWO kernels side by side: four warps per cloud for the
 * clouds with >= 336 distinct points (the critical path: the shortest round), and for the others a kernel that treats
 * cloud_weight[c] as the cloud's PERIOD (slot k = point k mod weight, what roipoint pooling's wrap produces) and spends a
 * round only on the distinct 32-slot blocks; the period is verified on the device and ignored when it does not hold, so
 * results never depend on it.  The second kernel runs on a stream owned by the library (one per device, created on
 * first use), forked from and joined to `stream` by events - legal under stream capture. */
MGAR_API size_t mgar_fps_ball_query_workspace_bytes(int b);
MGAR_API int mgar_fps_ball_query(int b, int n, int m, const float *xyz, int xyz_stride, float radius, int nsample,
                                 int *fps_idx, float *new_xyz, int *ball_idx, int *live_cnt, int write_dup_rows,
                                 const int *cloud_weight, void *workspace, size_t workspace_bytes, void *stream);

/* group_points_wrapper_stack: group_points.cpp:50-67 (kernel group_points_gpu.cu:71-102).
 * out (M,C,nsample), every element written. */
MGAR_API int mgar_group_points_stack(int B, int M, int C, int nsample, const float *features,
                            const int *features_batch_cnt, const int *idx, const int *idx_batch_cnt,
                            float *out, void *workspace, size_t workspace_bytes, void *stream);

/* group_points_grad_wrapper_stack: group_points.cpp:29-47 (kernel group_points_gpu.cu:15-44).
 * grad_features (N,C) zero-filled by the caller, accumulated with atomics. */
MGAR_API int mgar_group_points_grad_stack(int B, int M, int C, int N, int nsample, const float *grad_out,
                                 const int *idx, const int *idx_batch_cnt,
                                 const int *features_batch_cnt, float *grad_features, void *workspace, size_t workspace_bytes, void *stream);

/* ---- pointnet2_batch: channel-major gather / scatter, three-nearest interpolation ------------------- */

/* pointnet2_batch/src/pointnet2_api.cpp:11-23.  ball_query_wrapper of that module = mgar_ball_query_stack with
 * uniform_n = n, uniform_m = m, fill = 2; farthest_point_sampling_wrapper = mgar_fps_batch (same kernel source);
 * gather_points(_grad)_wrapper = the two entries below with nsample = 1.
 * group_points_wrapper_fast (group_points.cpp:39-54, kernel group_points_gpu.cu:55-72): points (b,c,n),
 * idx (b,npoints,nsample) -> out (b,c,npoints,nsample), every element written.  _grad (group_points.cpp:21-36,
 * kernel :13-31): grad_out (b,c,npoints,nsample) accumulated with atomics into zero-filled grad_points (b,c,n). */
MGAR_API int mgar_batch_group_points(int b, int c, int n, int npoints, int nsample, const float *points, const int *idx,
                                     float *out, void *stream);
MGAR_API int mgar_batch_group_points_grad(int b, int c, int n, int npoints, int nsample, const float *grad_out,
                                          const int *idx, float *grad_points, void *stream);
/* three_nn_wrapper_fast (interpolate.cpp:20-34, kernel interpolate_gpu.cu:14-62): unknown (b,n,3), known (b,m,3) ->
 * dist2 (b,n,3) SQUARED distances ascending, idx (b,n,3); ties keep the lowest index; with m < 3 the missing
 * slots are (+inf, 0).  three_interpolate_wrapper_fast (interpolate.cpp:37-54, kernel :82-101): points (b,c,m),
 * idx / weight (b,n,3) -> out (b,c,n); _grad (interpolate.cpp:57-73, kernel :124-146) accumulates into
 * zero-filled grad_points (b,c,m). */
MGAR_API int mgar_three_nn(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                           void *stream);
MGAR_API int mgar_three_interpolate(int b, int c, int m, int n, const float *points, const int *idx, const float *weight,
                                    float *out, void *stream);
MGAR_API int mgar_three_interpolate_grad(int b, int c, int n, int m, const float *grad_out, const int *idx,
                                         const float *weight, float *grad_points, void *stream);

/* Stack (ragged) layout of the same three ops.  three_nn_wrapper (pointnet2_stack/src/interpolate.cpp:32-60, kernel
 * interpolate_gpu.cu:16-74): unknown (N,3) / unknown_batch_cnt (B), known (M,3) / known_batch_cnt (B) -> dist2 (N,3)
 * SQUARED distances ascending, idx (N,3) GLOBAL known rows (cloud start added, :70-72); ties keep the lowest index;
 * a cloud with fewer than three known points leaves (+inf, cloud start).  three_interpolate_wrapper
 * (interpolate.cpp:63-83, kernel :107-126): features (M,C), idx / weight (N,3) -> out (N,C), every element written.
 * three_interpolate_grad_wrapper (interpolate.cpp:86-107, kernel :150-170): grad_out (N,C) accumulated with atomics
 * into zero-filled grad_features (M,C). */
MGAR_API int mgar_three_nn_stack(int B, int N, int M, const float *unknown, const int *unknown_batch_cnt,
                                 const float *known, const int *known_batch_cnt, float *dist2, int *idx, void *workspace, size_t workspace_bytes, void *stream);
MGAR_API int mgar_three_interpolate_stack(int N, int C, const float *features, const int *idx, const float *weight,
                                          float *out, void *stream);
MGAR_API int mgar_three_interpolate_grad_stack(int N, int C, const float *grad_out, const int *idx, const float *weight,
                                               float *grad_features, void *stream);

/* ---- Distance Aware Fusion Module: everything between the library GEMMs ------------------------- */

/* FusionAttention_mat.forward, model/gat_model.py:485-489 (RGB stream) and :501-505 (LiDAR stream), for B frames of n
 * persons and n_streams (1 or 2) streams in one launch:
 *   out[s] (B,n,D) = LayerNorm(softmax_keys((q[s] k[s]^T) * E / sqrt(D)) v[s] + resid[s]) * gamma[s] + beta[s].
 * q/k/v[s]: (B*n) rows of D floats with row stride ld (they may be column blocks of one projection GEMM output);
 * resid[s], out[s] (B,n,D) dense; E (B,n,n) = softmax(-De/sigma) (:484); valid (B,n) bytes, 0 = padded person whose
 * key is removed from the softmax (NULL: all valid); the pointer arrays are HOST arrays of n_streams DEVICE pointers.
 * n <= 128, D <= 512. */
MGAR_API int mgar_dafm_attention(int B, int n, int D, int n_streams, int ld, const float *const *q, const float *const *k,
                                 const float *const *v, const float *const *resid, const float *E,
                                 const unsigned char *valid, const float *const *gamma, const float *const *beta,
                                 float eps, float *const *out, void *stream);
/* out (rows,D) = LayerNorm(x + y) * gamma + beta (:490-491 / :506-507 after the FFN GEMMs); rows whose valid byte is 0
 * are written as zeros (the reference never computes them: it slices each frame to its person count, :1396-1405). */
MGAR_API int mgar_dafm_add_layernorm(int rows, int D, const float *x, const float *y, const float *gamma, const float *beta,
                                     const unsigned char *valid, float eps, float *out, void *stream);

/* ---- set-abstraction shared MLP (QueryAndGroup + [Conv1x1,BN,ReLU] x L + pool, fused) ------------ */

/* StackSAModuleMSG.forward for ONE scale, eval-mode BatchNorm: pointnet2_stack/pointnet2_modules.py:78-112
 * with QueryAndGroup (pointnet2_utils.py:123-159) fused in.  Inputs as for ball query plus
 * features (sum n_i, C_in) (NULL when C_in = 0) and idx (M,nsample) from ball query (rows whose
 * idx[0] < 0, or whose empty_mask byte is set, are empty balls: all-zero inputs, as
 * pointnet2_utils.py:146-150).  Ragged layout: pass both count arrays and uniform_n = 0.  Uniform
 * layout (every cloud uniform_n points and uniform_m centroids): count arrays may be NULL.
 * channels_host[0..n_layers] with channels_host[0] = 3 + C_in (HOST array).  params (DEVICE):
 * per layer W^T [c_in][c_out] then scale[c_out], shift[c_out] (BatchNorm folded:
 * scale = gamma/sqrt(var+eps), shift = beta - mean*scale).  out (M, C_last) with row stride
 * ld_out floats (so several scales can be written side by side).  pool_avg = 0: max, 1: mean.
 * Exact fp32 (FFMA) arithmetic. */
MGAR_API int mgar_sa_mlp_fwd_fp32(int B, int M, int nsample, int C_in, const float *xyz, const int *xyz_batch_cnt,
                                  const float *new_xyz, const int *new_xyz_batch_cnt, int uniform_n,
                                  int uniform_m, const float *features, const int *idx,
                                  const unsigned char *empty_mask, int n_layers, const int *channels_host,
                                  const float *params, float *out, int ld_out, int pool_avg, void *workspace, size_t workspace_bytes, void *stream);

/* Same operation on the 5th-generation tensor cores (tcgen05.mma kind::tf32, fp32 accumulation in
 * TMEM, 3xTF32 operand splitting: fp32-level accuracy, within the 1e-5 parity bar) for the shape
 * mgar_sa_mlp_tc_supported() accepts: nsample = 32, C_in <= 4, channels [3+C_in, 64, 128], max-pool
 * (BASELINE configs[1]).  Other shapes must use mgar_sa_mlp_fwd_fp32.  Extra arguments:
 * xyz_stride / feat_stride = floats between consecutive rows (3 / C_in for packed arrays; 4 / 4 with
 * features == xyz + 3 reads a pooled (N,3+1) array in place with one 128-bit load per point);
 * params_host = HOST copy of `params` (layer 1 is passed to the kernel by value);
 * out may be NULL when only the group maxima are wanted;
 * group_max (ceil(M/group_size), 128), optional: max of the output rows over each run of group_size
 * consecutive centroids (the per-person feature when every person has group_size centroids); rows
 * whose group_zero[g] != 0 stay zero (empty persons).  The buffer is cleared by this call.  * Empty-person skip: with out == NULL (ld_out is then otherwise unused) ld_out == -1 asks the kernel to walk only the
 * groups whose group_zero flag is 0 (a device-side compacted list; uniform layout with uniform_m == group_size,
 * group_size % 4 == 0, C_in = 1); flagged groups keep their zero row either way, so results are identical and
 * empty or zero-padded person slots cost nothing (one extra 1-CTA launch of ~4 us builds the list).
 * Distinct-centroid walk: group_live_cnt (ceil(M/group_size)) int32, optional, as written by mgar_fps_ball_query:
 * with the skip above in force the kernel walks, per live group, only the tiles of four centroids that cover its first
 * group_live_cnt[g] centroids - the others repeat centroid 0 (same ball, same feature row) and cannot change the
 * group's max, so group_max holds the same bits.
 * Distinct-sample walk: group_point_cnt (ceil(M/group_size)) int32, optional (nsample 32, with group_live_cnt):
 * group_point_cnt[g] = p states that group g's cloud holds p DISTINCT points and slot k repeats point k mod p - what
 * RoIPointPool3d's wrap produces (roipoint_pool3d_kernel.cu:86-97; min(in-box points, S), exported by
 * mgar_roipoint_pool_from_masks).  A ball row of such a cloud repeats its hits once it runs out of distinct ones (the
 * copies with index >= p, then the first hit as padding, ball_query_gpu.cu:49-60); the kernel then computes only the
 * octets of sample slots that start with a distinct sample, packed sixteen to a tile.  group_max holds the same maxima
 * to fp32 rounding of the operand split (<= 1e-6 relative, not the same bits: the power-of-two operand scale is shared
 * by different rows); a count that is not the cloud's period makes the result wrong - pass NULL when in doubt.
 * workspace: device scratch of mgar_sa_mlp_tc_workspace_bytes(M, group_size) bytes when the skip is asked for, of
 * mgar_stack_workspace_bytes(B) bytes with ragged batch counts (the two never combine: the skip needs the uniform
 * layout), NULL / 0 otherwise; the library allocates nothing. */
MGAR_API size_t mgar_sa_mlp_tc_workspace_bytes(int M, int group_size);
/* sample slots per list entry of the distinct-sample walk (8: a tile is 16 octets of one person) - for callers that want
 * to predict how many tiles a batch executes (bench.py's executed_fraction) */
MGAR_API int mgar_sa_mlp_tc_unit_slots(void);
MGAR_API int mgar_sa_mlp_tc_supported(int nsample, int C_in, int n_layers, const int *channels_host, int pool_avg);
/* nsample 16 and 64 also run on the tensor cores when ONLY group_max is wanted (out == NULL, uniform layout, whole
 * 32-position blocks per group): the kernel walks blocks of 32 positions - two 16-sample centroids or half a 64-sample
 * one - and the max over blocks equals the max over centroids.  This query answers for that call shape. */
MGAR_API int mgar_sa_mlp_tc_group_max_supported(int nsample, int C_in, int n_layers, const int *channels_host, int pool_avg);
MGAR_API int mgar_sa_mlp_fwd_tc(int B, int M, int nsample, int C_in, const float *xyz, int xyz_stride,
                                const int *xyz_batch_cnt, const float *new_xyz, const int *new_xyz_batch_cnt,
                                int uniform_n, int uniform_m, const float *features, int feat_stride, const int *idx,
                                const unsigned char *empty_mask, int n_layers, const int *channels_host,
                                const float *params, const float *params_host, float *out, int ld_out,
                                float *group_max, int group_size, const int *group_zero, const int *group_live_cnt,
                                const int *group_point_cnt, int pool_avg, void *workspace, size_t workspace_bytes,
                                void *stream);

/* ---- NeighborVoxelSAModuleMSG, one scale, eval mode (voxel_pool_modules.py:70-130) ------------------------------ */

/* out (M,C2) = ReLU(bn_out(W_out . pool_s ReLU(bn_in(W_in . features)[idx[m,s]] + bn_pos(W_pos . (xyz[idx[m,s]] - new_xyz[m])))))
 * fused: no (M,C,nsample) tensor is materialised.  features (N,C_in), xyz (N,3), new_xyz (M,3), idx (M,nsample) GLOBAL
 * rows as VoxelQuery returns them (rows of empty queries zeroed), empty_mask (M) uint8 (non-zero: the query found no
 * voxel; its grouped features and offsets count as zeros, as the reference masks them), params = the three layers with
 * BatchNorm folded, mgar_voxel_sa_param_floats(C_in, C1, C2) floats laid out
 *   W_in^T [C_in][C1] | scale_in | shift_in | W_pos^T [3][C1] | scale_pos | shift_pos | W_out^T [C1][C2] | scale_out | shift_out.
 * pool_avg: 0 = max over the samples, 1 = mean.  workspace: mgar_voxel_sa_workspace_bytes(N, C1) bytes (the (N,C1)
 * output of the input layer).  Channels <= 128.  Exact fp32 FFMA (the widths are 32-wide: a CUDA-core job). */
MGAR_API int mgar_voxel_sa_supported(int C_in, int C1, int C2, int nsample);
MGAR_API size_t mgar_voxel_sa_param_floats(int C_in, int C1, int C2);
MGAR_API size_t mgar_voxel_sa_workspace_bytes(int N, int C1);
MGAR_API int mgar_voxel_sa_fwd(int N, int M, int C_in, int C1, int C2, int nsample, int pool_avg, const float *features,
                               const float *xyz, const float *new_xyz, const int *idx, const unsigned char *empty_mask,
                               const float *params, float *out, int ld_out, void *workspace, size_t workspace_bytes,
                               void *stream);
/* mgar_voxel_sa_fwd for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels and same results as mgar_voxel_sa_fwd, without the workspace: the (N,C1)
 * scratch is a file-scope device array of MGAR_VOXEL_SA_STATIC_MAX_FLOATS floats (8 MiB per GPU that loads the library,
 * held until exit).  N * C1 beyond that is rejected with MGAR_ENOMEM.  The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine); its first call on a device resolves the array's device address
 * (cudaGetSymbolAddress) - make one call before starting a graph capture.  Every call on a device shares the array:
 * nothing in it has to start at zero (rows [0,N) are overwritten before they are read), but calls that overlap on the
 * GPU (two streams without a dependency) race on it - serialise them. */
#define MGAR_VOXEL_SA_STATIC_MAX_FLOATS (2u << 20)
MGAR_API int mgar_voxel_sa_fwd_static(int N, int M, int C_in, int C1, int C2, int nsample, int pool_avg,
                                      const float *features, const float *xyz, const float *new_xyz, const int *idx,
                                      const unsigned char *empty_mask, const float *params, float *out, int ld_out,
                                      void *stream);

/* ---- fp32-faithful GEMM on tcgen05 (3xTF32) --------------------------------------------------------------- */

/* C (M,N) = act(A (M,K) . Bt (N,K)^T + bias): the dense linears around the hot path that the reference runs through
 * torch (cuBLAS fp32): FusionAttention_mat's WQ/WK/WV products and FFN linears (model/gat_model.py:466-511), the
 * 128 -> 512 LiDAR embedding (:883-894).  A row-major (lda floats per row), Bt = the weight as nn.Linear stores it
 * ((out, in) row-major: K-major), bias (N) or NULL, relu != 0 applies max(., 0).  fp32 in / out; every product is
 * evaluated as lo.hi + hi.lo + hi.hi of tf32 halves with fp32 accumulation (error ~2^-22 relative per product).
 * Shapes: N % 128 == 0, K % 32 == 0 (mgar_gemm_tf32x3_supported); operands 16-byte aligned, ld* % 4 == 0. */
MGAR_API int mgar_gemm_tf32x3_supported(int M, int N, int K);
MGAR_API int mgar_gemm_tf32x3(int M, int N, int K, const float *A, int lda, const float *Bt, int ldb, const float *bias,
                              int relu, float *C, int ldc, void *stream);
/* The same product with the weight operand prepared ONCE: mgar_gemm_pack_bt splits Bt into tf32 hi / lo and stores every
 * (128 x 32) tile in the shared-memory image the MMA reads (mgar_gemm_packed_bytes(N, K) = 2 * N * K * 4 bytes, caller-
 * owned); mgar_gemm_tf32x3_packed then fetches a K chunk's weight tiles with one cp.async.bulk and its producer warps
 * only stream A.  Same results as mgar_gemm_tf32x3 bit for bit (same split, same accumulation order). */
MGAR_API size_t mgar_gemm_packed_bytes(int N, int K);
MGAR_API int mgar_gemm_pack_bt(int N, int K, const float *Bt, int ldb, float *packed, void *stream);
MGAR_API int mgar_gemm_tf32x3_packed(int M, int N, int K, const float *A, int lda, const float *packed, const float *bias,
                                     int relu, float *C, int ldc, void *stream);
/* `groups` (1 .. 4) independent products of ONE shape in ONE launch (host arrays of per-group operands; bias may be NULL, or
 * hold NULL entries): their tiles fill the SMs together - a (6400 x 512) x 512 product alone is 200 tiles on 148 SMs, two
 * waves for 1.35 waves of work.  The linears of FusionAttention_mat come in such pairs (the two streams; gat_model.py:466-511). */
MGAR_API int mgar_gemm_tf32x3_packed_grouped(int groups, int M, int N, int K, const float *const *A, const int *lda,
                                             const float *const *packed, const float *const *bias, int relu,
                                             float *const *C, const int *ldc, void *stream);

/* ---- training: the same module with BatchNorm batch statistics, and its backward pass ------------------- */

/* StackSAModuleMSG.forward in train() for ONE scale (pointnet2_modules.py:78-112; BatchNorm2d in training mode:
 * statistics over all M*nsample positions, biased variance, eps) and the backward pass that autograd derives for
 * it in the reference (cuDNN / ATen kernels + GroupingOperation.backward, group_points_gpu.cu:15-44).
 * weights (DEVICE): per layer W^T [c_in][c_out], concatenated.  gamma_beta (DEVICE): per layer gamma[c_out] then
 * beta[c_out].  batch_stats (optional): per layer mean[c_out] then BIASED variance[c_out] - what the caller needs
 * to update running_mean / running_var.  workspace: mgar_sa_mlp_train_workspace_bytes() bytes of device memory;
 * the forward leaves the activations there and mgar_sa_mlp_bwd must be given the SAME, untouched workspace.
 * Backward outputs: grad_weights / grad_gamma_beta (same layouts, overwritten) and grad_features (N, C_in), which
 * is ACCUMULATED into (zero it first; may be NULL).  No gradient is produced for xyz / new_xyz.
 * Geometry arguments as for mgar_sa_mlp_fwd_fp32.  Exact fp32 arithmetic, fp64 batch reductions. */
MGAR_API size_t mgar_sa_mlp_train_workspace_bytes(int M, int nsample, int n_layers, const int *channels_host);
/* The same given the pooling method (what mgar_sa_mlp_fwd_train / _bwd need at least for that call): layers
 * [3+C_in, 64, 128] with C_in 1 or 2, nsample 32 and max-pool - the hot-path shape - run specialised kernels that
 * keep no (positions x channels) activation, so their workspace is ~30x smaller.  Never larger than the above. */
MGAR_API size_t mgar_sa_mlp_train_workspace_bytes_for(int M, int nsample, int n_layers, const int *channels_host,
                                                      int pool_avg);
MGAR_API int mgar_sa_mlp_fwd_train(int B, int M, int nsample, int C_in, const float *xyz, const int *xyz_batch_cnt,
                                   const float *new_xyz, const int *new_xyz_batch_cnt, int uniform_n, int uniform_m,
                                   const float *features, const int *idx, const unsigned char *empty_mask,
                                   int n_layers, const int *channels_host, const float *weights,
                                   const float *gamma_beta, float eps, float *out, int ld_out, int pool_avg,
                                   float *batch_stats, void *workspace, size_t ws_bytes, void *stream);
MGAR_API int mgar_sa_mlp_bwd(int B, int M, int nsample, int C_in, const int *xyz_batch_cnt, const int *new_xyz_batch_cnt,
                             int uniform_n, int uniform_m, const int *idx, const unsigned char *empty_mask, int n_layers,
                             const int *channels_host, const float *weights, const float *gamma_beta,
                             const float *grad_out, int ld_grad, int pool_avg, void *workspace, size_t ws_bytes,
                             float *grad_weights, float *grad_gamma_beta, float *grad_features, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MGAR_LIDAR_H */
