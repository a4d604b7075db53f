/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C) of the reference algorithms on the
 * LiDAR hot path of kdu3613/Multimodal_GAR (vendored OpenPCDet ops).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call this;
 * the product (multimodal_gar_b200/csrc) never links or loads it.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (see oracle/Makefile).
 * -ffp-contract=off matters: every fused multiply-add the reference's compiled GPU code
 * performs is written as an explicit fmaf() here, everything else must stay unfused.
 *
 * Parity pinning: the reference ships no tests / golden vectors, so this
 * oracle is pinned against the reference itself: (i) orc_points_in_boxes_mask against the
 * reference's points_in_boxes_cpu compiled unmodified into oracle/_ref (tests/test_oracle_pin.py,
 * runs on CPU), (ii) every other function against the reference CUDA extensions in oracle/_ref on
 * the GPU box (tests/test_ref_gpu_parity.py, -m gpu), and the committed fixtures in tests/golden/.
 *
 * All citations are relative to /root/reference/pcdet/ops/.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---- in-box test -------------------------------------------------------------------------
 * roiaware_pool3d/src/roiaware_pool3d_kernel.cu:16-36 (GPU, MARGIN 1e-5; identical copy at
 * roipoint_pool3d/src/roipoint_pool3d_kernel.cu:15-35) and roiaware_pool3d/src/roiaware_pool3d.cpp:121-140
 * (CPU twin, MARGIN 1e-2).  gpu_contract=1 reproduces the FMA contraction nvcc emits for the
 * GPU kernels (checked in SASS: lx = fma(sx,cos, sy*(-sin)), ly = fma(sy,cos, sx*sin));
 * gpu_contract=0 is the host build (x86-64 baseline, no FMA contraction).
 * cosa/sina are passed in so callers can perturb them (ambiguity analysis below). */
static int in_box_cs(const float *pt, const float *box, float margin_f, int gpu_contract,
                     float cosa, float sina)
{
    float x = pt[0], y = pt[1], z = pt[2];
    float cx = box[0], cy = box[1], cz = box[2];
    float dx = box[3], dy = box[4], dz = box[5];
    if ((double)fabsf(z - cz) > (double)dz / 2.0) return 0;
    float sx = x - cx, sy = y - cy;
    float lx, ly;
    if (gpu_contract) {
        lx = fmaf(sx, cosa, sy * (-sina));
        ly = fmaf(sy, cosa, sx * sina);
    } else {
        float a = sx * cosa, b = sy * (-sina);
        lx = a + b;
        float c = sx * sina, d = sy * cosa;
        ly = c + d;
    }
    double m = (double)margin_f;
    int fx = (double)fabsf(lx) < (double)dx / 2.0 + m;
    int fy = (double)fabsf(ly) < (double)dy / 2.0 + m;
    return fx & fy;
}

static int in_box(const float *pt, const float *box, float margin_f, int gpu_contract)
{
    float rz = box[6];
    return in_box_cs(pt, box, margin_f, gpu_contract, cosf(-rz), sinf(-rz));
}

/* 1 when the in-box decision could flip if cosf/sinf differed by one ulp (glibc here versus
 * libdevice on the GPU): such (point, box) pairs are reported, not compared, by the CPU-side
 * parity tests; the GPU-vs-reference-GPU tests compare every pair. */
static int in_box_ambiguous(const float *pt, const float *box, float margin_f, int gpu_contract)
{
    float rz = box[6];
    float c = cosf(-rz), s = sinf(-rz);
    int base = in_box_cs(pt, box, margin_f, gpu_contract, c, s);
    float cs[3] = { nextafterf(c, -2.0f), c, nextafterf(c, 2.0f) };
    float ss[3] = { nextafterf(s, -2.0f), s, nextafterf(s, 2.0f) };
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            if (in_box_cs(pt, box, margin_f, gpu_contract, cs[i], ss[j]) != base) return 1;
    return 0;
}

/* points_in_boxes_gpu semantics: roiaware_pool3d_kernel.cu:313-336.  out (B,N) is pre-filled
 * with -1 by the caller (roiaware_pool3d_utils.py:38); only hits are written. */
void orc_points_in_boxes_first(const float *boxes, const float *pts, int *out, int B, int M, int N)
{
    const float margin = 1e-5f;
    for (int b = 0; b < B; b++)
        for (int i = 0; i < N; i++) {
            const float *p = pts + ((size_t)b * N + i) * 3;
            for (int k = 0; k < M; k++)
                if (in_box(p, boxes + ((size_t)b * M + k) * 7, margin, 1)) {
                    out[(size_t)b * N + i] = k;
                    break;
                }
        }
}

/* points_in_boxes_cpu semantics: roiaware_pool3d.cpp:143-168. out (M,N) 0/1, margin 1e-2. */
void orc_points_in_boxes_mask(const float *boxes, const float *pts, int *out, int M, int N)
{
    const float margin = 1e-2f;
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++)
            out[(size_t)i * N + j] = in_box(pts + (size_t)j * 3, boxes + (size_t)i * 7, margin, 0);
}

/* (M,N) 0/1 mask with the GPU kernel's semantics (margin 1e-5, FMA contraction):
 * roipoint_pool3d_kernel.cu:38-60 transposed to (M,N). */
void orc_in_box_mask_gpu_sem(const float *boxes, const float *pts, int *out, int M, int N)
{
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++)
            out[(size_t)i * N + j] = in_box(pts + (size_t)j * 3, boxes + (size_t)i * 7, 1e-5f, 1);
}

/* ambiguity map, same layout; cpu_sem selects margin 1e-2 / no contraction. */
void orc_in_box_ambiguous(const float *boxes, const float *pts, int *out, int M, int N, int cpu_sem)
{
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++)
            out[(size_t)i * N + j] = in_box_ambiguous(pts + (size_t)j * 3, boxes + (size_t)i * 7,
                                                      cpu_sem ? 1e-2f : 1e-5f, cpu_sem ? 0 : 1);
}

/* ---- roipoint_pool3d ---------------------------------------------------------------------
 * roipoint_pool3d_kernel.cu:38-134 (assign_pts_to_box3d, get_pooled_idx, roipool3d_forward).
 * boxes are the already-enlarged boxes (enlarge_box3d is applied by the Python wrapper,
 * roipoint_pool3d_utils.py:48).  pooled (B,M,S,3+C) and empty (B,M) are zero-filled by the caller.
 * pts_idx (B,M,S) may be NULL; when given it receives the pooled point indices (scratch in the
 * reference, exposed here for tests). */
void orc_roipoint_pool(const float *xyz, const float *boxes, const float *feat, float *pooled,
                       int *empty, int *pts_idx, int B, int N, int M, int S, int C)
{
    int *idx = (int *)malloc(sizeof(int) * (size_t)(S > 0 ? S : 1));
    for (int b = 0; b < B; b++)
        for (int m = 0; m < M; m++) {
            const float *box = boxes + ((size_t)b * M + m) * 7;
            int cnt = 0;
            for (int k = 0; k < N; k++) {
                if (in_box(xyz + ((size_t)b * N + k) * 3, box, 1e-5f, 1)) {
                    if (cnt < S) idx[cnt++] = k;
                    else break;
                }
            }
            size_t bm = (size_t)b * M + m;
            if (cnt == 0) {
                empty[bm] = 1;
                continue;           /* pooled stays zero; pts_idx row is uninitialised in the reference */
            }
            for (int k = cnt; k < S; k++) idx[k] = idx[k % cnt];
            for (int s = 0; s < S; s++) {
                int src = idx[s];
                float *dst = pooled + (bm * S + s) * (size_t)(3 + C);
                for (int j = 0; j < 3; j++) dst[j] = xyz[((size_t)b * N + src) * 3 + j];
                for (int j = 0; j < C; j++) dst[3 + j] = feat[((size_t)b * N + src) * C + j];
                if (pts_idx) pts_idx[bm * S + s] = src;
            }
        }
    free(idx);
}

/* ---- farthest point sampling ---------------------------------------------------------------
 * pointnet2/pointnet2_stack/src/sampling_gpu.cu:9-13 (opt_n_threads), :16-21 (__update),
 * :25-140 (batch kernel), :188-319 (stack kernel, block fixed at 1024, :340).
 * The block is simulated literally: thread tid scans k = tid, tid+bs, ... keeping the first
 * strict maximum, then the shared-memory tree halves with strict '>' (ties keep the lower slot).
 * Squared distance as nvcc 12.9 compiles it for sm_100 (SASS of oracle/_ref, every instantiation,
 * also ball_query_kernel_stack): FMUL on dy, then FFMA dx, then FFMA dz, i.e.
 * fma(dz,dz, fma(dx,dx, dy*dy)), d = x2 - x1. */
int orc_opt_n_threads(int work_size)
{
    int pow_2 = (int)(log((double)work_size) / log(2.0));
    int v = 1 << pow_2;
    if (v > 1024) v = 1024;
    if (v < 1) v = 1;
    return v;
}

static void fps_one(const float *xyz, float *temp, int n, int m, int bs, int *idxs, int idx_base)
{
    if (m <= 0) return;
    float *dists = (float *)malloc(sizeof(float) * bs);
    int *dists_i = (int *)malloc(sizeof(int) * bs);
    int old = 0;
    idxs[0] = idx_base;
    for (int j = 1; j < m; j++) {
        float x1 = xyz[old * 3 + 0], y1 = xyz[old * 3 + 1], z1 = xyz[old * 3 + 2];
        for (int tid = 0; tid < bs; tid++) {
            int besti = 0;
            float best = -1.0f;
            for (int k = tid; k < n; k += bs) {
                float dx = xyz[k * 3 + 0] - x1, dy = xyz[k * 3 + 1] - y1, dz = xyz[k * 3 + 2] - z1;
                float d = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
                float d2 = fminf(d, temp[k]);
                temp[k] = d2;
                if (d2 > best) { besti = k; best = d2; }
            }
            dists[tid] = best;
            dists_i[tid] = besti;
        }
        for (int half = bs / 2; half >= 1; half /= 2)
            for (int tid = 0; tid < half; tid++) {
                float v1 = dists[tid], v2 = dists[tid + half];
                int i1 = dists_i[tid], i2 = dists_i[tid + half];
                dists[tid] = v1 > v2 ? v1 : (v2 > v1 ? v2 : v1);   /* max(v1, v2) */
                dists_i[tid] = v2 > v1 ? i2 : i1;
            }
        old = dists_i[0];
        idxs[j] = old + idx_base;
    }
    free(dists);
    free(dists_i);
}

/* batch layout: xyz (b,n,3), temp (b,n) pre-filled 1e10 by the caller (pointnet2_utils.py:178),
 * idx (b,m) local indices. */
void orc_fps_batch(int b, int n, int m, const float *xyz, float *temp, int *idx)
{
    int bs = orc_opt_n_threads(n);
    for (int i = 0; i < b; i++)
        fps_one(xyz + (size_t)i * n * 3, temp + (size_t)i * n, n, m, bs, idx + (size_t)i * m, 0);
}

/* stack layout: xyz (sum n_i,3), idx (sum m_i) GLOBAL row indices (sampling_gpu.cu:225,317). */
void orc_fps_stack(int batch, const float *xyz, float *temp, const int *xyz_batch_cnt, int *idx,
                   const int *num_sampled)
{
    size_t start = 0, istart = 0;
    for (int i = 0; i < batch; i++) {
        fps_one(xyz + start * 3, temp + start, xyz_batch_cnt[i], num_sampled[i], 1024, idx + istart,
                (int)start);
        start += xyz_batch_cnt[i];
        istart += num_sampled[i];
    }
}

/* ---- ball query (stack) ------------------------------------------------------------------
 * pointnet2/pointnet2_stack/src/ball_query_gpu.cu:16-66.  idx (M,nsample) is zero-filled by
 * the caller (pointnet2_utils.py:33); slots the kernel does not write keep the caller's value. */
void orc_ball_query_stack(int B, int M, float radius, int nsample, const float *new_xyz,
                          const int *new_xyz_batch_cnt, const float *xyz, const int *xyz_batch_cnt,
                          int *idx)
{
    float radius2 = radius * radius;
    for (int pt = 0; pt < M; pt++) {
        int bs_idx = 0, pt_cnt = new_xyz_batch_cnt[0];
        for (int k = 1; k < B; k++) {
            if (pt < pt_cnt) break;
            pt_cnt += new_xyz_batch_cnt[k];
            bs_idx = k;
        }
        size_t start = 0;
        for (int k = 0; k < bs_idx; k++) start += xyz_batch_cnt[k];
        const float *q = new_xyz + (size_t)pt * 3;
        const float *p = xyz + start * 3;
        int *o = idx + (size_t)pt * nsample;
        int n = xyz_batch_cnt[bs_idx], cnt = 0;
        for (int k = 0; k < n; k++) {
            float dx = q[0] - p[k * 3 + 0], dy = q[1] - p[k * 3 + 1], dz = q[2] - p[k * 3 + 2];
            float d2 = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
            if (d2 < radius2) {
                if (cnt == 0)
                    for (int l = 0; l < nsample; l++) o[l] = k;
                o[cnt] = k;
                if (++cnt >= nsample) break;
            }
        }
        if (cnt == 0) o[0] = -1;
    }
}

/* ---- grouping (stack) ----------------------------------------------------------------------
 * pointnet2/pointnet2_stack/src/group_points_gpu.cu:71-102 (forward), :15-44 (grad).
 * out (M,C,nsample); idx holds batch-local indices. */
static size_t batch_of(int pt, int B, const int *idx_batch_cnt, const int *features_batch_cnt)
{
    int bs_idx = 0, pt_cnt = idx_batch_cnt[0];
    for (int k = 1; k < B; k++) {
        if (pt < pt_cnt) break;
        pt_cnt += idx_batch_cnt[k];
        bs_idx = k;
    }
    size_t start = 0;
    for (int k = 0; k < bs_idx; k++) start += features_batch_cnt[k];
    return start;
}

void orc_group_points_stack(int B, int M, int C, int nsample, const float *features,
                            const int *features_batch_cnt, const int *idx, const int *idx_batch_cnt,
                            float *out)
{
    for (int pt = 0; pt < M; pt++) {
        size_t start = batch_of(pt, B, idx_batch_cnt, features_batch_cnt);
        for (int c = 0; c < C; c++)
            for (int s = 0; s < nsample; s++)
                out[((size_t)pt * C + c) * nsample + s] =
                    features[(start + idx[(size_t)pt * nsample + s]) * C + c];
    }
}

/* grad_features (N,C) zero-filled by the caller (pointnet2_utils.py:101); the reference uses
 * float atomicAdd (order-nondeterministic); this restatement sums in (pt, c, s) order. */
void orc_group_points_grad_stack(int B, int M, int C, int N, int nsample, const float *grad_out,
                                 const int *idx, const int *idx_batch_cnt,
                                 const int *features_batch_cnt, float *grad_features)
{
    (void)N;
    for (int pt = 0; pt < M; pt++) {
        size_t start = batch_of(pt, B, idx_batch_cnt, features_batch_cnt);
        for (int c = 0; c < C; c++)
            for (int s = 0; s < nsample; s++)
                grad_features[(start + idx[(size_t)pt * nsample + s]) * C + c] +=
                    grad_out[((size_t)pt * C + c) * nsample + s];
    }
}

/* ---- RoI-aware voxel pooling ----------------------------------------------------
 * roiaware_pool3d/src/roiaware_pool3d_kernel.cu:39-198 (forward) and :236-312 (backward).
 * voxel index per axis (:62-68): int((local + d/2) / (d/out)) through `unsigned` with min/max against ints:
 * negative ints wrap and land on out-1.  The CUDA float->int cast saturates and maps NaN to 0; restated here. */
static unsigned voxel_axis(float local, float d, int out)
{
    float res = d / (float)out;
    float v = (local + d / 2) / res;
    int i;
    if (v != v) i = 0;
    else if (v >= 2147483648.0f) i = 2147483647;
    else if (v <= -2147483648.0f) i = (-2147483647 - 1);
    else i = (int)v;
    unsigned u = (unsigned)i;
    unsigned hi = (unsigned)(out - 1);
    return u < hi ? u : hi;
}

/* lists (boxes,ox,oy,oz,max_pts) zero-filled by the caller; argmax / pooled as the reference leaves them
 * (pooled only written for non-empty voxels).  pool_method 0 = max, 1 = avg. */
void orc_roiaware_pool3d(const float *rois, const float *pts, const float *feat, int *argmax, int *lists,
                         float *pooled, int boxes_num, int pts_num, int C, int max_pts, int ox, int oy, int oz,
                         int pool_method)
{
    unsigned max_num = (unsigned)(max_pts - 1);
    for (int b = 0; b < boxes_num; b++) {
        const float *box = rois + (size_t)b * 7;
        float cosa = cosf(-box[6]), sina = sinf(-box[6]);
        int *vox = lists + (size_t)b * ox * oy * oz * max_pts;
        for (int k = 0; k < pts_num; k++) {                        /* :88-103, point-index order */
            const float *p = pts + (size_t)k * 3;
            if (!in_box_cs(p, box, 1e-5f, 1, cosa, sina)) continue;
            float sx = p[0] - box[0], sy = p[1] - box[1];
            float lx = fmaf(sx, cosa, sy * (-sina)), ly = fmaf(sy, cosa, sx * sina), lz = p[2] - box[2];
            unsigned xi = voxel_axis(lx, box[3], ox), yi = voxel_axis(ly, box[4], oy), zi = voxel_axis(lz, box[5], oz);
            size_t base = (((size_t)xi * oy + yi) * oz + zi) * max_pts;
            unsigned cnt = (unsigned)vox[base];
            if (cnt < max_num) {
                vox[base + cnt + 1] = k;
                vox[base]++;
            }
        }
        for (size_t v = 0; v < (size_t)ox * oy * oz; v++) {
            const int *l = vox + v * max_pts;
            size_t o = ((size_t)b * ox * oy * oz + v) * C;
            for (int c = 0; c < C; c++) {
                if (pool_method == 0) {                            /* :136-152 */
                    int arg = -1;
                    float best = -INFINITY;
                    for (int k = 1; k <= l[0]; k++) {
                        float f = feat[(size_t)l[k] * C + c];
                        if (f > best) { best = f; arg = l[k]; }
                    }
                    if (arg != -1) pooled[o + c] = best;
                    argmax[o + c] = arg;
                } else {                                           /* :182-191 */
                    float sum = 0;
                    for (int k = 1; k <= l[0]; k++) sum += feat[(size_t)l[k] * C + c];
                    if (l[0] > 0) pooled[o + c] = sum / l[0];
                }
            }
        }
    }
}

/* grad_in (pts_num,C) zero-filled by the caller; the reference accumulates with float atomics (order
 * nondeterministic), this restatement in (box, voxel, channel, k) order. */
void orc_roiaware_pool3d_backward(const int *lists, const int *argmax, const float *grad_out, float *grad_in,
                                  int boxes_num, int ox, int oy, int oz, int C, int max_pts, int pool_method)
{
    size_t voxels = (size_t)boxes_num * ox * oy * oz;
    for (size_t v = 0; v < voxels; v++)
        for (int c = 0; c < C; c++) {
            if (pool_method == 0) {                                /* :254-256 */
                int a = argmax[v * C + c];
                if (a != -1) grad_in[(size_t)a * C + c] += grad_out[v * C + c] * 1;
            } else {                                               /* :279-283 */
                const int *l = lists + v * max_pts;
                float cur = 1 / fmaxf((float)l[0], 1.0f);
                for (int k = 1; k <= l[0]; k++) grad_in[(size_t)l[k] * C + c] += grad_out[v * C + c] * cur;
            }
        }
}

/* ---- voxel query ------------------------------------------------------------------
 * pointnet2_stack/src/voxel_query_gpu.cu:10-89.  idx (M,nsample) zero-filled by the caller
 * (voxel_query_utils.py:33).  Distance as the kernel's SASS computes it (same contraction as ball query):
 * fma(dz,dz, fma(dx,dx, dy*dy)) with d = point - query; a candidate is skipped when d2 > radius*radius. */
void orc_voxel_query_stack(int M, int R1, int R2, int R3, int nsample, float radius, int z_range, int y_range,
                           int x_range, const float *new_xyz, const float *xyz, const int *new_coords,
                           const int *point_indices, int *idx)
{
    float radius2 = radius * radius;
    for (int q = 0; q < M; q++) {
        const float *c = new_xyz + (size_t)q * 3;
        const int *v = new_coords + (size_t)q * 4;
        int *row = idx + (size_t)q * nsample;
        int cnt = 0;
        for (int dz = -z_range; dz <= z_range; dz++) {
            int z = v[1] + dz;
            if (z < 0 || z >= R1) continue;
            for (int dy = -y_range; dy <= y_range; dy++) {
                int y = v[2] + dy;
                if (y < 0 || y >= R2) continue;
                for (int dx = -x_range; dx <= x_range; dx++) {
                    int x = v[3] + dx;
                    if (x < 0 || x >= R3) continue;
                    int nb = point_indices[(((size_t)v[0] * R1 + z) * R2 + y) * R3 + x];
                    if (nb < 0) continue;
                    float ex = xyz[(size_t)nb * 3] - c[0], ey = xyz[(size_t)nb * 3 + 1] - c[1],
                          ez = xyz[(size_t)nb * 3 + 2] - c[2];
                    float d2 = fmaf(ez, ez, fmaf(ex, ex, ey * ey));
                    if (d2 > radius2) continue;
                    if (cnt < nsample) {
                        if (cnt == 0)
                            for (int l = 0; l < nsample; l++) row[l] = nb;
                        row[cnt++] = nb;
                    }
                }
            }
        }
        if (cnt == 0) row[0] = -1;
    }
}

/* ---- pointnet2_batch surface -------------------------------------------------------
 * ball_query_kernel_fast, pointnet2_batch/src/ball_query_gpu.cu:15-51: like the stack kernel but a centroid
 * without neighbours leaves its (zero-filled) row untouched.  SASS: d2 = fma(dz,dz, fma(dx,dx, dy*dy)). */
void orc_ball_query_batch(int b, int n, int m, float radius, int nsample, const float *new_xyz, const float *xyz,
                          int *idx)
{
    float radius2 = radius * radius;
    for (int bs = 0; bs < b; bs++)
        for (int pt = 0; pt < m; pt++) {
            const float *c = new_xyz + ((size_t)bs * m + pt) * 3;
            const float *p = xyz + (size_t)bs * n * 3;
            int *row = idx + ((size_t)bs * m + pt) * nsample;
            int cnt = 0;
            for (int k = 0; k < n; k++) {
                float ex = c[0] - p[k * 3], ey = c[1] - p[k * 3 + 1], ez = c[2] - p[k * 3 + 2];
                float d2 = fmaf(ez, ez, fmaf(ex, ex, ey * ey));
                if (d2 < radius2) {
                    if (cnt == 0)
                        for (int l = 0; l < nsample; l++) row[l] = k;
                    row[cnt++] = k;
                    if (cnt >= nsample) break;
                }
            }
        }
}

/* three_nn_kernel_fast, interpolate_gpu.cu:14-62: bests kept as doubles initialised to 1e40, float distance with
 * the same contraction; outputs are the squared distances. */
void orc_three_nn(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx)
{
    for (int bs = 0; bs < b; bs++)
        for (int pt = 0; pt < n; pt++) {
            const float *u = unknown + ((size_t)bs * n + pt) * 3;
            const float *kn = known + (size_t)bs * m * 3;
            double best1 = 1e40, best2 = 1e40, best3 = 1e40;
            int i1 = 0, i2 = 0, i3 = 0;
            for (int k = 0; k < m; k++) {
                float ex = u[0] - kn[k * 3], ey = u[1] - kn[k * 3 + 1], ez = u[2] - kn[k * 3 + 2];
                float d = fmaf(ez, ez, fmaf(ex, ex, ey * ey));
                if (d < best1) { best3 = best2; i3 = i2; best2 = best1; i2 = i1; best1 = d; i1 = k; }
                else if (d < best2) { best3 = best2; i3 = i2; best2 = d; i2 = k; }
                else if (d < best3) { best3 = d; i3 = k; }
            }
            float *o = dist2 + ((size_t)bs * n + pt) * 3;
            int *oi = idx + ((size_t)bs * n + pt) * 3;
            o[0] = (float)best1; o[1] = (float)best2; o[2] = (float)best3;
            oi[0] = i1; oi[1] = i2; oi[2] = i3;
        }
}

/* three_nn_kernel_stack, pointnet2_stack/src/interpolate_gpu.cu:16-74: ragged batches; the serial walk of one
 * cloud's known points with double bests (1e40) and strict '<'; indices leave as GLOBAL known rows (:70-72). */
void orc_three_nn_stack(int batch, const float *unknown, const int *unknown_cnt, const float *known, const int *known_cnt,
                        float *dist2, int *idx)
{
    size_t u0 = 0;
    int k0 = 0;
    for (int bs = 0; bs < batch; bs++) {
        const float *kn = known + (size_t)k0 * 3;
        const int m = known_cnt[bs];
        for (int pt = 0; pt < unknown_cnt[bs]; pt++) {
            const float *u = unknown + (u0 + pt) * 3;
            double best1 = 1e40, best2 = 1e40, best3 = 1e40;
            int i1 = 0, i2 = 0, i3 = 0;
            for (int k = 0; k < m; k++) {
                float ex = u[0] - kn[k * 3], ey = u[1] - kn[k * 3 + 1], ez = u[2] - kn[k * 3 + 2];
                float d = fmaf(ez, ez, fmaf(ex, ex, ey * ey));
                if (d < best1) { best3 = best2; i3 = i2; best2 = best1; i2 = i1; best1 = d; i1 = k; }
                else if (d < best2) { best3 = best2; i3 = i2; best2 = d; i2 = k; }
                else if (d < best3) { best3 = d; i3 = k; }
            }
            float *o = dist2 + (u0 + pt) * 3;
            int *oi = idx + (u0 + pt) * 3;
            o[0] = (float)best1; o[1] = (float)best2; o[2] = (float)best3;
            oi[0] = i1 + k0; oi[1] = i2 + k0; oi[2] = i3 + k0;
        }
        u0 += unknown_cnt[bs];
        k0 += m;
    }
}

/* three_interpolate_kernel_stack, pointnet2_stack/src/interpolate_gpu.cu:107-126: features (M,C) row-major ->
 * out (N,C); same contraction as the batch kernel (SASS of the reference build: FMUL w1*f1, FFMA w0*f0, FFMA w2*f2). */
void orc_three_interpolate_stack(int n, int c, const float *features, const int *idx, const float *weight, float *out)
{
    for (int pt = 0; pt < n; pt++) {
        const int *i = idx + (size_t)pt * 3;
        const float *w = weight + (size_t)pt * 3;
        for (int ch = 0; ch < c; ch++)
            out[(size_t)pt * c + ch] = fmaf(w[2], features[(size_t)i[2] * c + ch],
                                            fmaf(w[0], features[(size_t)i[0] * c + ch], w[1] * features[(size_t)i[1] * c + ch]));
    }
}

/* three_interpolate_kernel_fast, interpolate_gpu.cu:82-101; SASS (FMUL, FFMA, FFMA): fma(w2,p2, fma(w0,p0, w1*p1)). */
void orc_three_interpolate(int b, int c, int m, int n, const float *points, const int *idx, const float *weight,
                           float *out)
{
    for (int bs = 0; bs < b; bs++)
        for (int ch = 0; ch < c; ch++) {
            const float *p = points + ((size_t)bs * c + ch) * m;
            for (int pt = 0; pt < n; pt++) {
                const int *i = idx + ((size_t)bs * n + pt) * 3;
                const float *w = weight + ((size_t)bs * n + pt) * 3;
                out[((size_t)bs * c + ch) * n + pt] = fmaf(w[2], p[i[2]], fmaf(w[0], p[i[0]], w[1] * p[i[1]]));
            }
        }
}
