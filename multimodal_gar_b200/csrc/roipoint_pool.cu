// Per-person point pooling: the first S in-box points of every box, in point-index order.
//
// Reference: roipoint_pool3d/src/roipoint_pool3d_kernel.cu:38-134 - three kernels around a
// materialised (B,N,M) int matrix, the middle one (get_pooled_idx) a single thread per box walking
// N stride-M loads, plus two cudaMalloc/cudaFree per call.
// Here: ONE kernel, one CTA per (frame, box).  Each warp owns a contiguous slice of the frame's
// points and compacts its hits in order with ballot/popc into a private shared-memory list (no
// block barrier inside the scan, a conservative AABB vote skips the exact test for almost every
// warp-step); one barrier later the per-warp lists are concatenated (exclusive scan of 32 counts),
// the `k mod cnt` wrap is applied and xyz+features are gathered - optionally already in the
// box's canonical frame (pointrcnn_head.py:121-129).  No (N,M) intermediate, no allocation.
#include "mgar_common.cuh"

namespace mgar {

constexpr int kPoolUnroll = 4;

struct PoolBox {
    BoxP b;
    float R;  // conservative xy half-extent: a hit always has |x-cx| <= R and |y-cy| <= R
};

__global__ void __launch_bounds__(1024)
roipoint_pool_kernel(const float *__restrict__ xyz, const float *__restrict__ boxes,
                     const float *__restrict__ feat, float *__restrict__ pooled, int *__restrict__ empty,
                     int *__restrict__ pts_idx, int N, int M, int S, int C, int cap, int canonical) {
    extern __shared__ int smem_i[];
    const int nwarps = blockDim.x >> 5;
    int *wlist = smem_i;                  // [nwarps][cap]
    int *flist = wlist + nwarps * cap;    // [S]
    __shared__ int wcount[32];
    __shared__ int wstart[33];
    __shared__ PoolBox sb;

    const int m = blockIdx.x, b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t bm = (size_t)b * M + m;
    xyz += (size_t)b * N * 3;
    feat += (size_t)b * N * C;

    if (threadIdx.x == 0) {
        PoolBox p;
        p.b = make_boxp<false>(boxes + bm * 7, nullptr);
        // |(lx,ly)| equals |(sx,sy)| up to a few ulp, so hypot(tx,ty) with 0.1 % slack bounds |sx|,|sy|.
        p.R = sqrtf(p.b.tx * p.b.tx + p.b.ty * p.b.ty) * 1.001f;
        sb = p;
    }
    __syncthreads();
    const BoxP bp = sb.b;
    const float R = sb.R;

    // ---- phase 1: per-warp ordered compaction over a contiguous slice ------------------------
    const int seg = ((N + nwarps - 1) / nwarps + 31) & ~31;
    const int beg = min(N, warp * seg), end = min(N, beg + seg);
    int *mylist = wlist + warp * cap;
    int cnt = 0;
    for (int base = beg; base < end && cnt < S; base += 32 * kPoolUnroll) {
        float x[kPoolUnroll], y[kPoolUnroll], z[kPoolUnroll];
#pragma unroll
        for (int u = 0; u < kPoolUnroll; u++) {
            const int k = base + u * 32 + lane;
            if (k < end) {
                x[u] = __ldg(xyz + (size_t)k * 3 + 0);
                y[u] = __ldg(xyz + (size_t)k * 3 + 1);
                z[u] = __ldg(xyz + (size_t)k * 3 + 2);
            } else {
                x[u] = y[u] = z[u] = nanf("");
            }
        }
#pragma unroll
        for (int u = 0; u < kPoolUnroll; u++) {
            const int k = base + u * 32 + lane;
            const bool reject = !(k < end) || fabsf(__fsub_rn(z[u], bp.cz)) > bp.hz ||
                                fabsf(__fsub_rn(x[u], bp.cx)) > R || fabsf(__fsub_rn(y[u], bp.cy)) > R;
            if (__all_sync(0xffffffffu, reject)) continue;
            const bool hit = !reject && in_box<false>(x[u], y[u], z[u], bp);
            const unsigned msk = __ballot_sync(0xffffffffu, hit);
            if (msk == 0) continue;
            const int pos = cnt + __popc(msk & ((1u << lane) - 1u));
            if (hit && pos < cap) mylist[pos] = k;
            cnt += __popc(msk);
        }
    }
    if (lane == 0) wcount[warp] = min(cnt, cap);
    __syncthreads();

    // ---- phase 2: concatenate the per-warp lists in warp (= index) order ----------------------
    if (warp == 0) {
        int c = lane < nwarps ? wcount[lane] : 0;
        int incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        wstart[lane] = incl - c;
        if (lane == 31) wstart[32] = incl;
    }
    __syncthreads();
    const int total = min(wstart[32], S);
    {
        const int st = wstart[warp], c = wcount[warp];
        for (int i = lane; i < c && st + i < S; i += 32) flist[st + i] = mylist[i];
    }
    __syncthreads();

    // ---- phase 3: wrap (slot k >= cnt copies slot k mod cnt) + gather ---------------------------
    const int D = 3 + C;
    float *out = pooled + bm * (size_t)S * D;
    if (threadIdx.x == 0) empty[bm] = total == 0 ? 1 : 0;
    if (total == 0) {
        for (size_t i = threadIdx.x; i < (size_t)S * D; i += blockDim.x) out[i] = 0.f;
        if (pts_idx)
            for (int s = threadIdx.x; s < S; s += blockDim.x) pts_idx[bm * S + s] = 0;
        return;
    }
    const bool vec4 = (C == 1) && ((reinterpret_cast<uintptr_t>(pooled) & 15) == 0);
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        const int src = flist[s < total ? s : s % total];
        float px = __ldg(xyz + (size_t)src * 3 + 0);
        float py = __ldg(xyz + (size_t)src * 3 + 1);
        float pz = __ldg(xyz + (size_t)src * 3 + 2);
        if (canonical) {
            px = __fsub_rn(px, bp.cx); py = __fsub_rn(py, bp.cy); pz = __fsub_rn(pz, bp.cz);
            const float rx = __fmaf_rn(py, -bp.sina, __fmul_rn(px, bp.cosa));
            const float ry = __fmaf_rn(py, bp.cosa, __fmul_rn(px, bp.sina));
            px = rx; py = ry;
        }
        if (vec4) {
            reinterpret_cast<float4 *>(out)[s] = make_float4(px, py, pz, __ldg(feat + src));
        } else {
            float *o = out + (size_t)s * D;
            o[0] = px; o[1] = py; o[2] = pz;
            for (int j = 0; j < C; j++) o[3 + j] = __ldg(feat + (size_t)src * C + j);
        }
        if (pts_idx) pts_idx[bm * S + s] = src;
    }
}

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_roipoint_pool(const float *xyz, const float *boxes, const float *feat, float *pooled,
                                  int *empty, int *pts_idx, int B, int N, int M, int S, int C,
                                  int canonical, void *stream) {
    MGAR_REQUIRE(B >= 0 && N >= 0 && M >= 0 && S >= 0 && C >= 0, "roipoint_pool: negative size");
    if (B == 0 || M == 0) return MGAR_OK;
    MGAR_REQUIRE(pooled && empty && boxes && (xyz || N == 0) && (feat || C == 0 || N == 0), "roipoint_pool: null pointer");
    MGAR_REQUIRE(B <= 65535, "roipoint_pool: B=%d exceeds grid.y limit", B);
    // Largest warp count whose private lists + the final list fit in shared memory.
    int nwarps = 32, cap = 0;
    size_t smem = 0;
    for (; nwarps >= 1; nwarps >>= 1) {
        const int seg = ((N + nwarps - 1) / nwarps + 31) & ~31;
        cap = S < seg ? S : seg;
        if (cap < 1) cap = 1;
        smem = ((size_t)nwarps * cap + (size_t)(S > 0 ? S : 1)) * sizeof(int);
        if (smem <= 200 * 1024) break;
    }
    MGAR_REQUIRE(nwarps >= 1, "roipoint_pool: num_sampled_points=%d does not fit in shared memory", S);
    static bool attr_set = false;
    if (!attr_set) {
        MGAR_CUDA(cudaFuncSetAttribute(roipoint_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    dim3 grid(M, B);
    roipoint_pool_kernel<<<grid, nwarps * 32, smem, (cudaStream_t)stream>>>(xyz, boxes, feat, pooled, empty, pts_idx,
                                                                        N, M, S, C, cap, canonical);
    count_launch();
    return check_launch("roipoint_pool");
}
