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
#include <stdlib.h>

#include "mgar_common.cuh"

namespace mgar {

constexpr int kPoolUnroll = 4;

struct PoolBox {
    BoxP b;
    float R;  // conservative xy half-extent: a hit always has |x-cx| <= R and |y-cy| <= R
};

// wrap (slot k >= cnt copies slot k mod cnt, roipoint_pool3d_kernel.cu:86-97) + gather of xyz | features into the
// pooled rows, optionally in the box's canonical frame (pointrcnn_head.py:121-129).  flist: the first `total`
// in-box point indices in index order (shared memory).  Whole-CTA call.
__device__ __forceinline__ void pool_emit(const float *__restrict__ xyz, const float *__restrict__ feat,
                                          float *__restrict__ pooled, int *__restrict__ empty, int *__restrict__ pts_idx,
                                          const int *flist, int total, size_t bm, int S, int C, int canonical,
                                          const BoxP &bp, int *__restrict__ cnt = nullptr) {
    const int D = 3 + C;
    float *out = pooled + bm * (size_t)S * D;
    if (threadIdx.x == 0) {
        empty[bm] = total == 0 ? 1 : 0;
        if (cnt) cnt[bm] = total;              // distinct pooled points (<= S): slots beyond it repeat them
    }
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

    pool_emit(xyz, feat, pooled, empty, pts_idx, flist, total, bm, S, C, canonical, bp);
}

// ---- two-pass pooling over in-box BIT MASKS (the throughput path) -------------------------------------------
// Pass 1 (box_mask_kernel) streams the points ONCE for all boxes of the frame: every thread owns 4 consecutive
// points (three 128-bit loads), the pre-processed boxes are broadcast from shared memory, and for each box the
// 128 in-box bits of a warp's points are assembled with one REDUX.OR per 32-point word - written only when a
// word has a hit (the mask is zero-filled first; person boxes hold ~0.5 % of a frame).  The same tests give the
// first-hit index of points_in_boxes_gpu (roiaware_pool3d_kernel.cu:313-336) when the caller wants it.
// Pass 2 (pool_from_mask_kernel), one CTA per (frame, box), scans the box's N/32 words (8 KB instead of the
// frame's 786 KB), turns popcounts into ordered positions with one block scan, and emits the pooled rows.
// The (frame, box) CTA of roipoint_pool_kernel re-reads the whole frame per box: 40x the point traffic.
constexpr int kMaskThreads = 256;
constexpr int kMaskBoxTile = 512;

template <bool kSame>
__global__ void __launch_bounds__(kMaskThreads)
box_mask_kernel(const float *__restrict__ boxes_first, const float *__restrict__ boxes_pool, const float *__restrict__ pts,
                int *__restrict__ box_idx, unsigned *__restrict__ mask, int M, int N, int W, int vec_ok) {
    __shared__ BoxP sbox[kMaskBoxTile];
    __shared__ BoxP sfirst[kSame ? 1 : kMaskBoxTile];
    const int b = blockIdx.y, lane = threadIdx.x & 31;
    boxes_pool += (size_t)b * M * 7;
    if (boxes_first) boxes_first += (size_t)b * M * 7;
    pts += (size_t)b * N * 3;
    mask += (size_t)b * M * W;
    const long long p0 = ((long long)blockIdx.x * kMaskThreads + threadIdx.x) * 4;
    const int n_valid = p0 >= N ? 0 : (N - p0 >= 4 ? 4 : (int)(N - p0));
    float x[4], y[4], z[4];
    if (vec_ok && n_valid == 4) {
        const float4 *v = reinterpret_cast<const float4 *>(pts + p0 * 3);
        const float4 a = __ldg(v), bb = __ldg(v + 1), c = __ldg(v + 2);
        x[0] = a.x; y[0] = a.y; z[0] = a.z;   x[1] = a.w; y[1] = bb.x; z[1] = bb.y;
        x[2] = bb.z; y[2] = bb.w; z[2] = c.x; x[3] = c.y; y[3] = c.z; z[3] = c.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const bool ok = j < n_valid;
            x[j] = ok ? __ldg(pts + (p0 + j) * 3 + 0) : nanf("");   // NaN never tests in-box
            y[j] = ok ? __ldg(pts + (p0 + j) * 3 + 1) : nanf("");
            z[j] = ok ? __ldg(pts + (p0 + j) * 3 + 2) : nanf("");
        }
    }
    const int word0 = (int)((((long long)blockIdx.x * kMaskThreads + (threadIdx.x & ~31)) * 4) >> 5) + (lane >> 3);
    const unsigned sub = 0xffu << (lane & 24);          // the 8 lanes whose nibbles make one 32-point word
    const int sh = 4 * (lane & 7);
    int res[4] = {-1, -1, -1, -1};
    for (int k0 = 0; k0 < M; k0 += kMaskBoxTile) {
        const int kt = min(kMaskBoxTile, M - k0);
        __syncthreads();
        for (int k = threadIdx.x; k < kt; k += kMaskThreads) {
            sbox[k] = make_boxp<false>(boxes_pool + (size_t)(k0 + k) * 7, nullptr);
            if (!kSame) sfirst[k] = make_boxp<false>(boxes_first + (size_t)(k0 + k) * 7, nullptr);
        }
        __syncthreads();
        for (int k = 0; k < kt; k++) {
            const BoxP bp = sbox[k];
            unsigned nib = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const bool h = in_box<false>(x[j], y[j], z[j], bp);
                nib |= (h ? 1u : 0u) << j;
                if (kSame) { if (box_idx && h && res[j] < 0) res[j] = k0 + k; }
            }
            if (!kSame) {
                const BoxP bf = sfirst[k];
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (res[j] < 0 && in_box<false>(x[j], y[j], z[j], bf)) res[j] = k0 + k;
            }
            if (__any_sync(0xffffffffu, nib != 0)) {
                const unsigned word = __reduce_or_sync(sub, nib << sh);
                if ((lane & 7) == 0 && word != 0 && word0 < W) mask[(size_t)(k0 + k) * W + word0] = word;
            }
        }
    }
    if (box_idx) {
        box_idx += (size_t)b * N;
        if (vec_ok && n_valid == 4) {
            *reinterpret_cast<int4 *>(box_idx + p0) = make_int4(res[0], res[1], res[2], res[3]);
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (j < n_valid) box_idx[p0 + j] = res[j];
        }
    }
}

// Same outputs with a per-frame BOX GRID: person boxes are small (R ~ 0.6 m) in a scene tens of metres wide, so a
// point only needs the boxes whose conservative xy square covers its 1 m cell - typically none or one instead of
// all M.  Every CTA rebuilds the frame's grid in shared memory (64 x 64 cells of 1 m around the origin, border cells
// absorb everything outside; up to 8 box ids per cell) - M * ~9 shared atomics against 1024 points * M tests saved.
// A hit sets its mask bit with one atomicOr (hits are ~0.5 % of the pairs); the first-hit index is the minimum hit.
// Exactness: the cell index is a monotone function of x, and a box is entered into every cell between
// cell(cx - R') and cell(cx + R') with R' >= |x - cx| of any in-box point (hypot of the half extents + margin, 0.1 %
// slack), so an in-box point always finds its box; the in-box test itself is unchanged.  A cell that would overflow
// switches the CTA to the exhaustive loop (crowds of overlapping boxes).
constexpr int kGridN = 64, kGridCap = 8, kGridMaxBoxes = 255;
constexpr size_t kGridSmem = (size_t)kGridN * kGridN * (4 + kGridCap) + (size_t)(kGridMaxBoxes + 1) * sizeof(BoxP);
constexpr int kGridIters = 4;                              // point tiles (1024 points each) per CTA
constexpr float kGridOrigin = -32.f, kGridInv = 1.f;     // cell = floor((x - origin) * inv), clamped to [0, 64)

__device__ __forceinline__ int grid_cell(float x) {
    const float c = floorf(__fmul_rn(__fsub_rn(x, kGridOrigin), kGridInv));
    return (int)fminf(fmaxf(c, 0.f), (float)(kGridN - 1));          // NaN -> 0 (fmaxf drops it)
}

__global__ void __launch_bounds__(kMaskThreads)
box_mask_grid_kernel(const float *__restrict__ boxes, const float *__restrict__ pts, int *__restrict__ box_idx,
                     unsigned *__restrict__ mask, int M, int N, int W, int vec_ok) {
    extern __shared__ int grid_sm[];
    int *scnt = grid_sm;                                                              // [cells] boxes per cell
    unsigned char (*sid)[kGridCap] = reinterpret_cast<unsigned char (*)[kGridCap]>(scnt + kGridN * kGridN);   // [cells][cap] box ids
    BoxP *sbox = reinterpret_cast<BoxP *>(sid + kGridN * kGridN);                     // [M]
    __shared__ int s_overflow;
    const int b = blockIdx.y;
    boxes += (size_t)b * M * 7;
    pts += (size_t)b * N * 3;
    mask += (size_t)b * M * W;
    for (int i = threadIdx.x; i < kGridN * kGridN; i += kMaskThreads) scnt[i] = 0;
    if (threadIdx.x == 0) s_overflow = 0;
    __syncthreads();
    for (int k = threadIdx.x; k < M; k += kMaskThreads) {
        const BoxP bp = make_boxp<false>(boxes + (size_t)k * 7, nullptr);
        sbox[k] = bp;
        const float R = sqrtf(bp.tx * bp.tx + bp.ty * bp.ty) * 1.001f + 1e-6f;
        const int x0 = grid_cell(bp.cx - R), x1 = grid_cell(bp.cx + R), y0 = grid_cell(bp.cy - R), y1 = grid_cell(bp.cy + R);
        if (!(R == R) || !(bp.cx == bp.cx) || !(bp.cy == bp.cy) || (x1 - x0 + 1) * (y1 - y0 + 1) > 64) {
            s_overflow = 1;                                            // NaN box or a huge one: exhaustive loop
        } else {
            for (int cy = y0; cy <= y1; cy++)
                for (int cx = x0; cx <= x1; cx++) {
                    const int old = atomicAdd(&scnt[cy * kGridN + cx], 1);
                    if (old >= kGridCap) s_overflow = 1;
                    else sid[cy * kGridN + cx][old] = (unsigned char)k;
                }
        }
    }
    __syncthreads();
    for (int it = 0; it < kGridIters; it++) {   // several point tiles per CTA amortise the grid build
        const long long p0 = (((long long)blockIdx.x * kGridIters + it) * kMaskThreads + threadIdx.x) * 4;
        const int n_valid = p0 >= N ? 0 : (N - p0 >= 4 ? 4 : (int)(N - p0));
        float x[4], y[4], z[4];
        if (vec_ok && n_valid == 4) {
            const float4 *v = reinterpret_cast<const float4 *>(pts + p0 * 3);
            const float4 a = __ldg(v), bb = __ldg(v + 1), c = __ldg(v + 2);
            x[0] = a.x; y[0] = a.y; z[0] = a.z;   x[1] = a.w; y[1] = bb.x; z[1] = bb.y;
            x[2] = bb.z; y[2] = bb.w; z[2] = c.x; x[3] = c.y; y[3] = c.z; z[3] = c.w;
        } else {
    #pragma unroll
            for (int j = 0; j < 4; j++) {
                const bool ok = j < n_valid;
                x[j] = ok ? __ldg(pts + (p0 + j) * 3 + 0) : nanf("");
                y[j] = ok ? __ldg(pts + (p0 + j) * 3 + 1) : nanf("");
                z[j] = ok ? __ldg(pts + (p0 + j) * 3 + 2) : nanf("");
            }
        }
        int res[4] = {-1, -1, -1, -1};
        if (s_overflow) {
            for (int k = 0; k < M; k++) {
                const BoxP bp = sbox[k];
    #pragma unroll
                for (int j = 0; j < 4; j++)
                    if (in_box<false>(x[j], y[j], z[j], bp)) {
                        if (res[j] < 0) res[j] = k;
                        atomicOr(mask + (size_t)k * W + ((p0 + j) >> 5), 1u << ((p0 + j) & 31));
                    }
            }
        } else {
    #pragma unroll
            for (int j = 0; j < 4; j++) {
                const int c = grid_cell(y[j]) * kGridN + grid_cell(x[j]);
                const int cnt = scnt[c];
                for (int q = 0; q < cnt; q++) {
                    const int k = sid[c][q];
                    if (in_box<false>(x[j], y[j], z[j], sbox[k])) {
                        res[j] = res[j] < 0 ? k : min(res[j], k);
                        atomicOr(mask + (size_t)k * W + ((p0 + j) >> 5), 1u << ((p0 + j) & 31));
                    }
                }
            }
        }
        if (box_idx) {
            int *box_idx_f = box_idx + (size_t)b * N;
            if (vec_ok && n_valid == 4) {
                *reinterpret_cast<int4 *>(box_idx_f + p0) = make_int4(res[0], res[1], res[2], res[3]);
            } else {
    #pragma unroll
                for (int j = 0; j < 4; j++)
                    if (j < n_valid) box_idx_f[p0 + j] = res[j];
            }
        }
    }
}

// ---- the same pass with the box grid built ONCE PER FRAME -----------------------------------------------------------
// box_mask_grid_kernel rebuilds the 64 x 64 grid in each of a frame's 16 CTAs (shared-memory atomics, two barriers, 52 KB
// of shared memory -> 3 CTAs per SM, 31 % occupancy, 1.3 TB/s).  Here a one-CTA-per-frame kernel builds the frame's
// grid in caller scratch (mgar_box_masks_workspace_bytes: per frame 4096 cell counts, 4096 x 8 box ids, the
// pre-processed boxes, an overflow flag), and the streaming kernel only reads it: the cell of a point through the
// read-only path (a frame's 36 KB of cells stay in L1 / L2), the boxes from a 1.3 KB shared-memory copy.  No grid build,
// no barrier but the one after the box copy, ~2 KB of shared memory: the SM is full of warps with loads in flight.
struct FrameGrid {                      // layout of one frame's scratch (bytes)
    static __host__ __device__ size_t cnt_off() { return 0; }
    static __host__ __device__ size_t ids_off() { return (size_t)kGridN * kGridN; }
    static __host__ __device__ size_t flag_off() { return ids_off() + (size_t)kGridN * kGridN * kGridCap; }
    static __host__ __device__ size_t box_off() { return flag_off() + 16; }
    static __host__ __device__ size_t bytes(int M) { return (box_off() + (size_t)M * sizeof(BoxP) + 255) & ~(size_t)255; }
};

// grid (1 + kZeroCtas, B): CTA 0 of a frame builds its grid; the others zero-fill the frame's in-box bit masks (mask != null:
// the masks are OR-ed into afterwards) - one launch instead of a memset node + a 64-CTA kernel one after the other.
constexpr int kZeroCtas = 16;
__global__ void __launch_bounds__(kMaskThreads)
frame_grid_build_kernel(const float *__restrict__ boxes, unsigned char *__restrict__ ws, int M, unsigned *__restrict__ mask,
                        long long mask_words) {
    extern __shared__ int grid_sm[];
    int *scnt = grid_sm;                                                              // [cells]
    __shared__ int s_overflow;
    const int b = blockIdx.y;
    if (blockIdx.x > 0) {
        if (!mask) return;
        uint4 *dst = reinterpret_cast<uint4 *>(mask + (size_t)b * mask_words);          // (the launcher checked the alignment)
        const long long n16 = mask_words / 4;
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (long long i = (long long)(blockIdx.x - 1) * kMaskThreads + threadIdx.x; i < n16; i += (long long)kZeroCtas * kMaskThreads)
            dst[i] = z;
        return;
    }
    boxes += (size_t)b * M * 7;
    ws += (size_t)b * FrameGrid::bytes(M);
    unsigned char *gcnt = ws + FrameGrid::cnt_off();
    unsigned char *gids = ws + FrameGrid::ids_off();
    BoxP *gbox = reinterpret_cast<BoxP *>(ws + FrameGrid::box_off());
    for (int i = threadIdx.x; i < kGridN * kGridN; i += kMaskThreads) scnt[i] = 0;
    if (threadIdx.x == 0) s_overflow = 0;
    __syncthreads();
    for (int k = threadIdx.x; k < M; k += kMaskThreads) {
        const BoxP bp = make_boxp<false>(boxes + (size_t)k * 7, nullptr);
        gbox[k] = bp;
        const float R = sqrtf(bp.tx * bp.tx + bp.ty * bp.ty) * 1.001f + 1e-6f;
        const int x0 = grid_cell(bp.cx - R), x1 = grid_cell(bp.cx + R), y0 = grid_cell(bp.cy - R), y1 = grid_cell(bp.cy + R);
        if (!(R == R) || !(bp.cx == bp.cx) || !(bp.cy == bp.cy) || (x1 - x0 + 1) * (y1 - y0 + 1) > 64) {
            s_overflow = 1;                                            // NaN box or a huge one: exhaustive loop
        } else {
            for (int cy = y0; cy <= y1; cy++)
                for (int cx = x0; cx <= x1; cx++) {
                    // a cell with more than kGridCap boxes is marked (count 255): only the points that fall into it take
                    // the exhaustive loop.  Zero-padded proposal slots are degenerate boxes at the origin and all land in
                    // the same 2 x 2 cells - round 1 (and the shared-memory grid kernel) gave up the whole frame for them.
                    const int old = atomicAdd(&scnt[cy * kGridN + cx], 1);
                    if (old < kGridCap) gids[(size_t)(cy * kGridN + cx) * kGridCap + old] = (unsigned char)k;
                }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kGridN * kGridN; i += kMaskThreads) gcnt[i] = (unsigned char)(scnt[i] > kGridCap ? 255 : scnt[i]);
    if (threadIdx.x == 0) *reinterpret_cast<int *>(ws + FrameGrid::flag_off()) = s_overflow;
}

#ifndef MGAR_GRID_ITERS
#define MGAR_GRID_ITERS 2
#endif
constexpr int kGridLIters = MGAR_GRID_ITERS;                // point tiles (1024 points each) per CTA
constexpr int kTileBytes = kMaskThreads * 4 * 3 * (int)sizeof(float);   // one tile of xyz rows: 12 KB

// ---- bulk asynchronous copy (the TMA engine's 1-D form): a CTA's point tiles are in flight before it does anything else --
__device__ __forceinline__ uint32_t pm_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pm_bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void pm_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}

// kBulk: the CTA's kGridLIters tiles of 1024 points (12 KB each) are brought into shared memory by cp.async.bulk copies
// issued before anything else (one mbarrier per tile); a thread then reads its four points with three conflict-free
// 128-bit shared loads.  Needs 16-byte aligned rows and whole tiles (the launcher checks; partial tiles use the other
// instantiation's loads).
template <bool kBulk>
__global__ void __launch_bounds__(kMaskThreads)
box_mask_framegrid_kernel(const unsigned char *__restrict__ ws, const float *__restrict__ pts, int *__restrict__ box_idx,
                          unsigned *__restrict__ mask, int M, int N, int W, int vec_ok) {
    extern __shared__ __align__(16) int grid_sm[];
    __shared__ __align__(8) unsigned long long s_bar[kGridLIters];
    // [kBulk: kGridLIters point tiles] [cell counts: kGridN^2 bytes] [M boxes]
    unsigned char *scnt = reinterpret_cast<unsigned char *>(grid_sm) + (kBulk ? kGridLIters * kTileBytes : 0);
    BoxP *sbox = reinterpret_cast<BoxP *>(scnt + kGridN * kGridN);
    const int b = blockIdx.y;
    ws += (size_t)b * FrameGrid::bytes(M);
    const unsigned char *gcnt = ws + FrameGrid::cnt_off();
    const uint2 *gids = reinterpret_cast<const uint2 *>(ws + FrameGrid::ids_off());  // 8 ids of a cell = one 8-byte load
    const int overflow = __ldg(reinterpret_cast<const int *>(ws + FrameGrid::flag_off()));
    pts += (size_t)b * N * 3;
    mask += (size_t)b * M * W;
    if (kBulk) {
        if (threadIdx.x == 0) {
#pragma unroll
            for (int it = 0; it < kGridLIters; it++)
                asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pm_smem_u32(&s_bar[it])), "r"(1) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
            for (int it = 0; it < kGridLIters; it++) {
                const long long t0 = ((long long)blockIdx.x * kGridLIters + it) * kMaskThreads * 4;       // first point of the tile
                if (t0 < N)
                    pm_bulk_load(pm_smem_u32(reinterpret_cast<unsigned char *>(grid_sm) + it * kTileBytes), pts + t0 * 3, kTileBytes,
                                 pm_smem_u32(&s_bar[it]));
            }
        }
    }
    {
        const float4 *src = reinterpret_cast<const float4 *>(ws + FrameGrid::box_off());
        float4 *dst = reinterpret_cast<float4 *>(sbox);
        for (int i = threadIdx.x; i < M * 2; i += kMaskThreads) dst[i] = __ldg(src + i);
        // the frame's cell counts (4 KB) into shared memory: a point's lookup is then a shared byte load instead of a
        // dependent global one (the kernel's top stall was the long scoreboard on exactly these loads)
        const uint4 *csrc = reinterpret_cast<const uint4 *>(gcnt);
        uint4 *cdst = reinterpret_cast<uint4 *>(scnt);
        for (int i = threadIdx.x; i < kGridN * kGridN / 16; i += kMaskThreads) cdst[i] = __ldg(csrc + i);
    }
    __syncthreads();
    for (int it = 0; it < kGridLIters; it++) {
        const long long p0 = (((long long)blockIdx.x * kGridLIters + it) * kMaskThreads + threadIdx.x) * 4;
        const int n_valid = p0 >= N ? 0 : (N - p0 >= 4 ? 4 : (int)(N - p0));
        float x[4], y[4], z[4];
        if (kBulk) {
            if (n_valid == 0) continue;                       // (whole tiles only: the tile is either complete or absent)
            pm_mbar_wait(pm_smem_u32(&s_bar[it]), 0);
            const float4 *v = reinterpret_cast<const float4 *>(reinterpret_cast<unsigned char *>(grid_sm) + it * kTileBytes) + threadIdx.x * 3;
            const float4 a = v[0], bb = v[1], c = v[2];
            x[0] = a.x; y[0] = a.y; z[0] = a.z;   x[1] = a.w; y[1] = bb.x; z[1] = bb.y;
            x[2] = bb.z; y[2] = bb.w; z[2] = c.x; x[3] = c.y; y[3] = c.z; z[3] = c.w;
        } else if (vec_ok && n_valid == 4) {
            const float4 *v = reinterpret_cast<const float4 *>(pts + p0 * 3);
            const float4 a = __ldg(v), bb = __ldg(v + 1), c = __ldg(v + 2);
            x[0] = a.x; y[0] = a.y; z[0] = a.z;   x[1] = a.w; y[1] = bb.x; z[1] = bb.y;
            x[2] = bb.z; y[2] = bb.w; z[2] = c.x; x[3] = c.y; y[3] = c.z; z[3] = c.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const bool ok = j < n_valid;
                x[j] = ok ? __ldg(pts + (p0 + j) * 3 + 0) : nanf("");
                y[j] = ok ? __ldg(pts + (p0 + j) * 3 + 1) : nanf("");
                z[j] = ok ? __ldg(pts + (p0 + j) * 3 + 2) : nanf("");
            }
        }
        int res[4] = {-1, -1, -1, -1};
        if (overflow) {
            for (int k = 0; k < M; k++) {
                const BoxP bp = sbox[k];
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (in_box<false>(x[j], y[j], z[j], bp)) {
                        if (res[j] < 0) res[j] = k;
                        atomicOr(mask + (size_t)k * W + ((p0 + j) >> 5), 1u << ((p0 + j) & 31));
                    }
            }
        } else {
            int cell[4], cnt[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {                     // the four cell counts in flight together
                cell[j] = grid_cell(y[j]) * kGridN + grid_cell(x[j]);
                cnt[j] = scnt[cell[j]];
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (cnt[j] == 0) continue;
                if (cnt[j] == 255) {                          // crowded cell: this point tests every box
                    for (int k = 0; k < M; k++)
                        if (in_box<false>(x[j], y[j], z[j], sbox[k])) {
                            if (res[j] < 0) res[j] = k;
                            atomicOr(mask + (size_t)k * W + ((p0 + j) >> 5), 1u << ((p0 + j) & 31));
                        }
                    continue;
                }
                const uint2 id8 = __ldg(gids + cell[j]);
                for (int q = 0; q < cnt[j]; q++) {
                    const int k = ((q < 4 ? id8.x : id8.y) >> (8 * (q & 3))) & 0xff;
                    if (in_box<false>(x[j], y[j], z[j], sbox[k])) {
                        res[j] = res[j] < 0 ? k : min(res[j], k);
                        atomicOr(mask + (size_t)k * W + ((p0 + j) >> 5), 1u << ((p0 + j) & 31));
                    }
                }
            }
        }
        if (box_idx) {
            int *box_idx_f = box_idx + (size_t)b * N;
            if (vec_ok && n_valid == 4) {
                *reinterpret_cast<int4 *>(box_idx_f + p0) = make_int4(res[0], res[1], res[2], res[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (j < n_valid) box_idx_f[p0 + j] = res[j];
            }
        }
    }
}

__global__ void __launch_bounds__(256)
pool_from_mask_kernel(const float *__restrict__ xyz, const float *__restrict__ boxes, const float *__restrict__ feat,
                      const unsigned *__restrict__ mask, float *__restrict__ pooled, int *__restrict__ empty,
                      int *__restrict__ pts_idx, int *__restrict__ cnt_out, int N, int M, int W, int S, int C, int canonical) {
    extern __shared__ int smem_i[];
    int *flist = smem_i;                  // [S]
    __shared__ int wsum[8];
    __shared__ int s_total;
    __shared__ BoxP sb;
    const int m = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t bm = (size_t)b * M + m;
    xyz += (size_t)b * N * 3;
    feat += (size_t)b * N * C;
    mask += bm * W;
    if (tid == 0) sb = make_boxp<false>(boxes + bm * 7, nullptr);
    // contiguous word range per thread -> ordered positions from one block-wide exclusive scan of popcounts
    const int per = (W + 255) / 256;
    const int w0 = min(W, tid * per), w1 = min(W, w0 + per);
    int cnt = 0;
    for (int w = w0; w < w1; w++) cnt += __popc(__ldg(mask + w));
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < warp; w++) base += wsum[w];
    if (tid == 255) s_total = base + incl;
    int pos = base + incl - cnt;
    for (int w = w0; w < w1 && pos < S; w++) {
        unsigned word = __ldg(mask + w);
        while (word && pos < S) {
            const int bit = __ffs(word) - 1;
            flist[pos++] = w * 32 + bit;
            word &= word - 1;
        }
    }
    __syncthreads();
    const int total = min(s_total, S);
    pool_emit(xyz, feat, pooled, empty, pts_idx, flist, total, bm, S, C, canonical, sb, cnt_out);
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
    // per call: the attribute belongs to the current device's context (a process may drive several GPUs)
    if (smem > 48 * 1024)
        MGAR_CUDA(cudaFuncSetAttribute(roipoint_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(M, B);
    roipoint_pool_kernel<<<grid, nwarps * 32, smem, (cudaStream_t)stream>>>(xyz, boxes, feat, pooled, empty, pts_idx,
                                                                        N, M, S, C, cap, canonical);
    count_launch();
    return check_launch("roipoint_pool");
}

extern "C" size_t mgar_box_mask_words(int N) { return (size_t)((N + 31) / 32); }

// scratch of the per-frame box grids (optional: without it every CTA builds its frame's grid in shared memory)
extern "C" size_t mgar_box_masks_workspace_bytes(int B, int M) {
    if (B <= 0 || M <= 0 || M > kGridMaxBoxes) return 0;
    return (size_t)B * FrameGrid::bytes(M);
}

extern "C" int mgar_box_masks(const float *boxes_first, const float *boxes_pool, const float *pts, int *box_idx,
                              unsigned *mask, int B, int M, int N, void *workspace, size_t workspace_bytes, void *stream) {
    MGAR_REQUIRE(B >= 0 && M >= 0 && N >= 0, "box_masks: negative size");
    if (B == 0 || N == 0 || M == 0) {
        if (box_idx && B > 0 && N > 0) MGAR_CUDA(cudaMemsetAsync(box_idx, 0xff, (size_t)B * N * 4, (cudaStream_t)stream));
        return MGAR_OK;
    }
    MGAR_REQUIRE(pts && boxes_pool && mask, "box_masks: null pointer");
    MGAR_REQUIRE(!box_idx || boxes_first, "box_masks: box_idx wanted but boxes_first is null");
    MGAR_REQUIRE(B <= 65535, "box_masks: B=%d exceeds grid.y limit", B);
    cudaStream_t st = (cudaStream_t)stream;
    const int W = (N + 31) / 32;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(pts) & 15) == 0) && (!box_idx || (reinterpret_cast<uintptr_t>(box_idx) & 15) == 0) &&
                       (N % 4 == 0);
    dim3 grid(div_up(N, kMaskThreads * 4), B);
    static_assert(sizeof(BoxP) == 32, "the frame grid copies boxes as two float4");
    const bool frame_grid = (!box_idx || boxes_first == boxes_pool) && M <= kGridMaxBoxes && !getenv("MGAR_NO_BOX_GRID") && workspace &&
        workspace_bytes >= mgar_box_masks_workspace_bytes(B, M) && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0 &&
        !getenv("MGAR_NO_FRAME_GRID");
    // the masks are cleared by the grid-building launch when a frame's words are whole 16-byte groups, by a memset otherwise
    const bool zero_in_kernel = frame_grid && ((size_t)M * W) % 4 == 0 && (reinterpret_cast<uintptr_t>(mask) & 15) == 0;
    if (!zero_in_kernel) MGAR_CUDA(cudaMemsetAsync(mask, 0, (size_t)B * M * W * sizeof(unsigned), st));
    if (frame_grid) {
        frame_grid_build_kernel<<<dim3(1 + (zero_in_kernel ? kZeroCtas : 0), B), kMaskThreads, kGridN * kGridN * sizeof(int), st>>>(
            boxes_pool, static_cast<unsigned char *>(workspace), M, zero_in_kernel ? mask : nullptr, (long long)M * W);
        count_launch();
        grid.x = div_up(N, kMaskThreads * 4 * kGridLIters);
        // MGAR_BOX_MASK_BULK=1 (whole 1024-point tiles of 16-byte aligned rows): the tiles are staged by the TMA engine's
        // bulk copies.  Measured on the bench step: 0.0525-0.0534 ms for the stage against 0.0515-0.0529 with per-thread
        // 128-bit loads - the kernel is bound by its ~4 instructions per point (cell arithmetic, byte load, test), not by
        // memory latency, so the deeper memory parallelism buys nothing; per-thread loads stay the default.
        const char *bulk_env = getenv("MGAR_BOX_MASK_BULK");
        const bool bulk = vec_ok && N % (kMaskThreads * 4) == 0 && ((size_t)N * 12) % 16 == 0 && bulk_env && bulk_env[0] == '1';
        if (bulk) {
            const size_t sm = (size_t)kGridLIters * kTileBytes + (size_t)kGridN * kGridN + (size_t)M * sizeof(BoxP);
            box_mask_framegrid_kernel<true><<<grid, kMaskThreads, sm, st>>>(static_cast<const unsigned char *>(workspace), pts, box_idx,
                                                                          mask, M, N, W, vec_ok);
        } else {
            box_mask_framegrid_kernel<false><<<grid, kMaskThreads, (size_t)kGridN * kGridN + (size_t)M * sizeof(BoxP), st>>>(
                static_cast<const unsigned char *>(workspace), pts, box_idx, mask, M, N, W, vec_ok);
        }
    }
    else if ((!box_idx || boxes_first == boxes_pool) && M <= kGridMaxBoxes && !getenv("MGAR_NO_BOX_GRID"))
    {
        MGAR_CUDA(cudaFuncSetAttribute(box_mask_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGridSmem));
        grid.x = div_up(N, kMaskThreads * 4 * kGridIters);
        box_mask_grid_kernel<<<grid, kMaskThreads, kGridSmem, st>>>(boxes_pool, pts, box_idx, mask, M, N, W, vec_ok);
    }
    else if (!box_idx || boxes_first == boxes_pool)
        box_mask_kernel<true><<<grid, kMaskThreads, 0, st>>>(boxes_first, boxes_pool, pts, box_idx, mask, M, N, W, vec_ok);
    else
        box_mask_kernel<false><<<grid, kMaskThreads, 0, st>>>(boxes_first, boxes_pool, pts, box_idx, mask, M, N, W, vec_ok);
    count_launch();
    return check_launch("box_masks");
}

extern "C" int mgar_roipoint_pool_from_masks(const float *xyz, const float *boxes, const float *feat, const unsigned *mask,
                                             float *pooled, int *empty, int *pts_idx, int *pooled_cnt, int B, int N, int M,
                                             int S, int C, int canonical, void *stream) {
    MGAR_REQUIRE(B >= 0 && N >= 0 && M >= 0 && S >= 0 && C >= 0, "roipoint_pool_from_masks: negative size");
    if (B == 0 || M == 0) return MGAR_OK;
    MGAR_REQUIRE(pooled && empty && boxes && mask && (xyz || N == 0) && (feat || C == 0 || N == 0), "roipoint_pool_from_masks: null pointer");
    MGAR_REQUIRE(B <= 65535, "roipoint_pool_from_masks: B=%d exceeds grid.y limit", B);
    const size_t smem = (size_t)(S > 0 ? S : 1) * sizeof(int);
    MGAR_REQUIRE(smem <= 200 * 1024, "roipoint_pool_from_masks: num_sampled_points=%d does not fit in shared memory", S);
    if (smem > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(pool_from_mask_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(M, B);
    pool_from_mask_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(xyz, boxes, feat, mask, pooled, empty, pts_idx, pooled_cnt, N, M,
                                                                     (N + 31) / 32, S, C, canonical);
    count_launch();
    return check_launch("roipoint_pool_from_masks");
}
