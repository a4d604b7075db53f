// RoI-aware voxel pooling: every box is cut into out_x*out_y*out_z voxels, the points inside a voxel
// are listed (first max_pts_each_voxel-1 in point-index order) and their features max- or average-pooled.
//
// Reference: roiaware_pool3d/src/roiaware_pool3d_kernel.cu:36-312 - a (boxes, npoints) int32 code matrix in a
// cudaMalloc'ed scratch, ONE THREAD PER BOX walking all npoints codes serially to build the voxel lists, then a
// pooling kernel with one thread per (voxel, channel) striding the feature rows by C floats.
// Here: pass 1 is the in-box bit-mask kernel the person pooling already uses (roipoint_pool.cu: one sweep over
// the points for all boxes, 1 bit per pair instead of 4 bytes); pass 2 gives every box a warp that visits only
// the non-zero mask words, recomputes the voxel code of the <= 32 hits of a word in parallel and appends them to
// their voxel lists in lane (= point-index) order with match_any/popc ranks - no serial walk, no code matrix;
// pass 3 pools with the channel as the fastest thread index, so a warp reads 128 contiguous bytes of one point.
// Results are bit-identical to the reference's (tests/test_roiaware_gpu.py runs the reference extension beside it).
#include "mgar_common.cuh"

namespace mgar {

constexpr int kRaWarps = 4;      // boxes per CTA in the collect pass
constexpr int kRaThreads = 256;  // pooling / backward CTA

// voxel index along one axis (roiaware_pool3d_kernel.cu:62-68): int((local + d/2) / (d/out)) with the reference's
// unsigned min/max: a negative int (or INT_MIN from -inf) wraps to a huge unsigned and lands on out-1.
__device__ __forceinline__ unsigned voxel_axis(float local, float d, int out) {
    const float res = __fdiv_rn(d, (float)out);
    const float v = __fdiv_rn(__fadd_rn(local, __fmul_rn(d, 0.5f)), res);
    const unsigned i = (unsigned)__float2int_rz(v);
    return min(i, (unsigned)(out - 1));
}

// One warp per box.  mask (boxes, W) in-box bits; lists (boxes, ox, oy, oz, max_pts) int32, slot 0 = count,
// zero-filled by the caller (the reference's contract, roiaware_pool3d_utils.py:71-72).
__global__ void __launch_bounds__(kRaWarps * 32)
roiaware_collect_kernel(const float *__restrict__ rois, const float *__restrict__ pts, const unsigned *__restrict__ mask,
                        int *__restrict__ lists, int boxes_num, int pts_num, int W, int max_pts, int ox, int oy, int oz) {
    const int lane = threadIdx.x & 31;
    const int box = blockIdx.x * kRaWarps + (threadIdx.x >> 5);
    if (box >= boxes_num) return;
    const float *r = rois + (size_t)box * 7;
    const BoxP b = make_boxp<false>(r, nullptr);
    const float dx = r[3], dy = r[4], dz = r[5];
    const unsigned *mrow = mask + (size_t)box * W;
    int *vox = lists + (size_t)box * ox * oy * oz * max_pts;
    const unsigned max_num = (unsigned)(max_pts - 1);
    const unsigned lt = (1u << lane) - 1u;

    for (int w0 = 0; w0 < W; w0 += 32) {
        const unsigned mw = (w0 + lane < W) ? mrow[w0 + lane] : 0u;
        unsigned nz = __ballot_sync(0xffffffffu, mw != 0u);
        while (nz) {
            const int j = __ffs(nz) - 1;
            nz &= nz - 1;
            const unsigned word = __shfl_sync(0xffffffffu, mw, j);
            if ((word >> lane) & 1u) {
                const int n = (w0 + j) * 32 + lane;
                const float x = pts[(size_t)n * 3], y = pts[(size_t)n * 3 + 1], z = pts[(size_t)n * 3 + 2];
                const float sx = __fsub_rn(x, b.cx), sy = __fsub_rn(y, b.cy);
                const float lx = __fmaf_rn(sx, b.cosa, __fmul_rn(sy, -b.sina));   // as in_box<false>
                const float ly = __fmaf_rn(sy, b.cosa, __fmul_rn(sx, b.sina));
                const float lz = __fsub_rn(z, b.cz);
                const unsigned xi = voxel_axis(lx, dx, ox), yi = voxel_axis(ly, dy, oy), zi = voxel_axis(lz, dz, oz);
                const size_t base = (((size_t)xi * oy + yi) * oz + zi) * max_pts;
                const unsigned peers = __match_any_sync(word, (xi << 16) | (yi << 8) | zi);
                const unsigned cnt = (unsigned)vox[base];
                const unsigned rank = __popc(peers & lt);
                __syncwarp(word);                                  // every peer has read the count
                if (cnt + rank < max_num) vox[base + cnt + rank + 1] = n;
                if (rank == 0 && cnt < max_num) vox[base] = (int)min(cnt + (unsigned)__popc(peers), max_num);
            }
            __syncwarp();                                          // counts written before the next word reads them
        }
    }
}

// thread = (box, voxel, channel), channel fastest.  pooled / argmax are written only where the reference writes
// them (non-empty voxels for pooled, roiaware_pool3d_kernel.cu:147-150, :189-191).
template <bool kMax>
__global__ void __launch_bounds__(kRaThreads)
roiaware_pool_kernel(const float *__restrict__ feat, const int *__restrict__ lists, float *__restrict__ pooled,
                     int *__restrict__ argmax, size_t total, int C, int max_pts) {
    const size_t t = (size_t)blockIdx.x * kRaThreads + threadIdx.x;
    if (t >= total) return;
    const size_t voxel = t / C;
    const int c = (int)(t - voxel * C);
    const int *l = lists + voxel * max_pts;
    const int cnt = l[0];
    if (kMax) {
        int arg = -1;
        float best = -INFINITY;                                    // float(-1e50)
        for (int k = 1; k <= cnt; ++k) {
            const int p = l[k];
            const float v = feat[(size_t)p * C + c];
            if (v > best) { best = v; arg = p; }
        }
        if (arg != -1) pooled[t] = best;
        argmax[t] = arg;
    } else {
        float sum = 0.f;
        for (int k = 1; k <= cnt; ++k) sum = __fadd_rn(sum, feat[(size_t)l[k] * C + c]);
        if (cnt > 0) pooled[t] = __fdiv_rn(sum, (float)cnt);
    }
}

template <bool kMax>
__global__ void __launch_bounds__(kRaThreads)
roiaware_pool_bwd_kernel(const int *__restrict__ lists, const int *__restrict__ argmax, const float *__restrict__ grad_out,
                         float *__restrict__ grad_in, size_t total, int C, int max_pts) {
    const size_t t = (size_t)blockIdx.x * kRaThreads + threadIdx.x;
    if (t >= total) return;
    const size_t voxel = t / C;
    const int c = (int)(t - voxel * C);
    if (kMax) {
        const int a = argmax[t];
        if (a == -1) return;
        atomicAdd(grad_in + (size_t)a * C + c, grad_out[t]);
    } else {
        const int *l = lists + voxel * max_pts;
        const int cnt = l[0];
        if (cnt <= 0) return;
        const float g = __fmul_rn(grad_out[t], __fdiv_rn(1.f, fmaxf((float)cnt, 1.f)));
        for (int k = 1; k <= cnt; ++k) atomicAdd(grad_in + (size_t)l[k] * C + c, g);
    }
}

}  // namespace mgar

using namespace mgar;

extern "C" size_t mgar_roiaware_pool3d_workspace_bytes(int boxes_num, int pts_num) {
    if (boxes_num <= 0 || pts_num <= 0) return 0;
    return (size_t)boxes_num * mgar_box_mask_words(pts_num) * sizeof(unsigned);
}

extern "C" int mgar_roiaware_pool3d_fwd(const float *rois, const float *pts, const float *pts_feature, int *argmax,
                                        int *pts_idx_of_voxels, float *pooled_features, void *workspace, int boxes_num,
                                        int pts_num, int channels, int max_pts_each_voxel, int out_x, int out_y,
                                        int out_z, int pool_method, void *stream) {
    MGAR_REQUIRE(boxes_num >= 0 && pts_num >= 0 && channels >= 0, "roiaware_pool3d: negative size");
    MGAR_REQUIRE(pool_method == 0 || pool_method == 1, "roiaware_pool3d: pool_method=%d (0 = max, 1 = avg)", pool_method);
    MGAR_REQUIRE(out_x >= 1 && out_y >= 1 && out_z >= 1 && out_x <= 256 && out_y <= 256 && out_z <= 256,
                 "roiaware_pool3d: out_size (%d,%d,%d) outside 1..256 (the voxel code holds 8 bits per axis)", out_x, out_y, out_z);
    MGAR_REQUIRE(max_pts_each_voxel >= 1, "roiaware_pool3d: max_pts_each_voxel=%d", max_pts_each_voxel);
    if (boxes_num == 0) return MGAR_OK;
    MGAR_REQUIRE(pts_idx_of_voxels && pooled_features && (argmax || pool_method == 1), "roiaware_pool3d: null output");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t voxels = (size_t)boxes_num * out_x * out_y * out_z;
    if (pts_num > 0) {
        MGAR_REQUIRE(rois && pts && workspace && (pts_feature || channels == 0), "roiaware_pool3d: null pointer");
        unsigned *mask = (unsigned *)workspace;
        const int rc = mgar_box_masks(nullptr, rois, pts, nullptr, mask, 1, boxes_num, pts_num, nullptr, 0, stream);
        if (rc != MGAR_OK) return rc;
        roiaware_collect_kernel<<<div_up(boxes_num, kRaWarps), kRaWarps * 32, 0, st>>>(
            rois, pts, mask, pts_idx_of_voxels, boxes_num, pts_num, (int)mgar_box_mask_words(pts_num), max_pts_each_voxel,
            out_x, out_y, out_z);
        count_launch();
    }
    const size_t total = voxels * channels;
    if (total > 0) {
        const size_t blocks = (total + kRaThreads - 1) / kRaThreads;
        MGAR_REQUIRE(blocks <= 0x7fffffffu, "roiaware_pool3d: %zu outputs exceed the grid", total);
        if (pool_method == 0)
            roiaware_pool_kernel<true><<<(unsigned)blocks, kRaThreads, 0, st>>>(pts_feature, pts_idx_of_voxels, pooled_features,
                                                                               argmax, total, channels, max_pts_each_voxel);
        else
            roiaware_pool_kernel<false><<<(unsigned)blocks, kRaThreads, 0, st>>>(pts_feature, pts_idx_of_voxels, pooled_features,
                                                                                argmax, total, channels, max_pts_each_voxel);
        count_launch();
    }
    return check_launch("roiaware_pool3d_fwd");
}

extern "C" int mgar_roiaware_pool3d_bwd(const int *pts_idx_of_voxels, const int *argmax, const float *grad_out,
                                        float *grad_in, int boxes_num, int out_x, int out_y, int out_z, int channels,
                                        int max_pts_each_voxel, int pool_method, void *stream) {
    MGAR_REQUIRE(boxes_num >= 0 && channels >= 0 && out_x >= 0 && out_y >= 0 && out_z >= 0, "roiaware_pool3d_backward: negative size");
    MGAR_REQUIRE(pool_method == 0 || pool_method == 1, "roiaware_pool3d_backward: pool_method=%d (0 = max, 1 = avg)", pool_method);
    const size_t total = (size_t)boxes_num * out_x * out_y * out_z * channels;
    if (total == 0) return MGAR_OK;
    MGAR_REQUIRE(grad_out && grad_in && (pool_method == 0 ? (const void *)argmax : (const void *)pts_idx_of_voxels),
                 "roiaware_pool3d_backward: null pointer");
    const size_t blocks = (total + kRaThreads - 1) / kRaThreads;
    MGAR_REQUIRE(blocks <= 0x7fffffffu, "roiaware_pool3d_backward: %zu outputs exceed the grid", total);
    cudaStream_t st = (cudaStream_t)stream;
    if (pool_method == 0)
        roiaware_pool_bwd_kernel<true><<<(unsigned)blocks, kRaThreads, 0, st>>>(pts_idx_of_voxels, argmax, grad_out, grad_in, total,
                                                                               channels, max_pts_each_voxel);
    else
        roiaware_pool_bwd_kernel<false><<<(unsigned)blocks, kRaThreads, 0, st>>>(pts_idx_of_voxels, argmax, grad_out, grad_in, total,
                                                                                channels, max_pts_each_voxel);
    count_launch();
    return check_launch("roiaware_pool3d_backward");
}
