// Farthest point sampling (batch and stack layouts), bit-exact with the reference's tie order.
//
// Reference: pointnet2/pointnet2_stack/src/sampling_gpu.cu:25-140 (batch) and :188-319 (stack):
// one CTA per cloud, every one of the m-1 rounds re-reads xyz and temp from global memory and
// runs a 10-level shared-memory tree with ~10 block barriers.
//
// Here the cloud and its running min-distances live in REGISTERS for the whole kernel; a round is
// a handful of FFMA per owned point, two warp REDUX (max of the distance bits, min of the
// tie-break rank) and - only when a cloud is spread over several warps - one block barrier.
//
// Tie order.  The reference's winner among equal distances is decided by its launch geometry:
// thread t scans k = t, t+bs, ... keeping the FIRST maximum (strict >), then the tree keeps the
// LOWER slot on ties at strides bs/2 ... 1.  Two tied threads first meet at the stride of their
// lowest differing bit and the one with that bit clear survives, i.e. the winner minimises
// (bitrev_L(k mod bs), k div bs) lexicographically, L = log2(bs).  bs = opt_n_threads(n) =
// min(1024, 2^floor(log2 n)) for the batch kernel (sampling_gpu.cu:9-13,150) and 1024 for the
// stack kernel (:340).  We enumerate the points in exactly that order ("rank" r) and hand rank r
// to thread r / PPL, slot r % PPL: the first strict maximum in (thread, slot) order is then the
// reference's winner, with no per-element tie handling in the hot loop.
#include <limits.h>
#include <stdlib.h>

#include "fps_bq_cta.cuh"
#include <mutex>

#include "fps_bq_dup.cuh"
#include "mgar_common.cuh"

namespace mgar {

struct FpsGeom {
    int L;    // log2(bs)
    int SH;   // log2(stripes padded to a power of two)
};

__device__ __forceinline__ int rank_to_index(int r, int L, int SH) {
    const int t = L == 0 ? 0 : (int)(__brev((unsigned)(r >> SH)) >> (32 - L));
    return ((r & ((1 << SH) - 1)) << L) + t;
}

// One CTA of W warps per cloud; each thread owns PPL ranks (registers).
template <int W, int PPL>
__global__ void __launch_bounds__(W * 32)
fps_reg_kernel(const float *__restrict__ xyz, int xs, float *__restrict__ temp, int *__restrict__ idxs,
               float *__restrict__ new_xyz, const int *__restrict__ xyz_batch_cnt,
               const int *__restrict__ num_sampled, int n_uniform, int m_uniform, int L_uniform, int stack,
               int smem_pts) {
    extern __shared__ float sxyz[];               // [n][3] staged cloud (lookup of the current point)
    __shared__ int s_red[2][W > 1 ? W : 1][2];    // double-buffered per-warp (value bits, rank)
    __shared__ int s_off[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bi = blockIdx.x;
    int n, m, L;
    long long start, istart;
    if (stack) {
        // exclusive prefix of the ragged counts (the reference loops serially: sampling_gpu.cu:207-211)
        int a = 0, c = 0;
        for (int k = tid; k < bi; k += W * 32) { a += xyz_batch_cnt[k]; c += num_sampled[k]; }
        a = __reduce_add_sync(0xffffffffu, a);
        c = __reduce_add_sync(0xffffffffu, c);
        if (tid == 0) { s_off[0] = 0; s_off[1] = 0; }
        __syncthreads();
        if (lane == 0) { atomicAdd(&s_off[0], a); atomicAdd(&s_off[1], c); }
        __syncthreads();
        start = s_off[0]; istart = s_off[1];
        n = xyz_batch_cnt[bi]; m = num_sampled[bi]; L = 10;
    } else {
        n = n_uniform; m = m_uniform; L = L_uniform;
        start = (long long)bi * n; istart = (long long)bi * m;
    }
    if (m <= 0) return;
    xyz += start * xs; idxs += istart;
    if (temp) temp += start;
    if (new_xyz) new_xyz += istart * 3;
    if (n > smem_pts || n < 1) {  // caller broke the max_n contract: make it visible, never overrun
        for (int j = tid; j < m; j += W * 32) idxs[j] = -1;
        return;
    }
    const int base = stack ? (int)start : 0;
    const int bs = 1 << L;
    int stripes = (n + bs - 1) >> L, SH = 0;
    while ((1 << SH) < stripes) SH++;

    if (xs == 3) {
        for (int i = tid; i < n * 3; i += W * 32) sxyz[i] = xyz[i];
    } else {   // strided rows (e.g. the xyz columns of a pooled (n,4) array)
        for (int i = tid; i < n * 3; i += W * 32) sxyz[i] = xyz[(size_t)(i / 3) * xs + (i % 3)];
    }
    __syncthreads();

    float px[PPL], py[PPL], pz[PPL], pt[PPL];
#pragma unroll
    for (int i = 0; i < PPL; i++) {
        const int k = rank_to_index(tid * PPL + i, L, SH);
        const bool valid = (tid * PPL + i) < (bs << SH) && k < n;
        px[i] = valid ? sxyz[k * 3 + 0] : 0.f;
        py[i] = valid ? sxyz[k * 3 + 1] : 0.f;
        pz[i] = valid ? sxyz[k * 3 + 2] : 0.f;
        pt[i] = valid ? (temp ? temp[k] : 1e10f) : -1.f;  // -1 never beats the reference's initial best = -1 either
    }

    int cur = 0, j = 1;
    if (tid == 0) idxs[0] = base;
    for (; j < m; j++) {
        const float cx = sxyz[cur * 3 + 0], cy = sxyz[cur * 3 + 1], cz = sxyz[cur * 3 + 2];
        float best = -1.f;
        int bslot = 0;
#pragma unroll
        for (int i = 0; i < PPL; i++) {
            const float d = sqdist(__fsub_rn(px[i], cx), __fsub_rn(py[i], cy), __fsub_rn(pz[i], cz));
            pt[i] = fminf(d, pt[i]);
            if (pt[i] > best) { best = pt[i]; bslot = i; }
        }
        // distances are >= 0 (or the -1 filler): their int bit patterns order like the floats
        const int vb = __float_as_int(best);
        int vmax = __reduce_max_sync(0xffffffffu, vb);
        int rmin = __reduce_min_sync(0xffffffffu, vb == vmax ? tid * PPL + bslot : INT_MAX);
        if (W > 1) {
            const int buf = j & 1;
            if (lane == 0) { s_red[buf][warp][0] = vmax; s_red[buf][warp][1] = rmin; }
            __syncthreads();
            if (W <= 8) {
                vmax = s_red[buf][0][0]; rmin = s_red[buf][0][1];
#pragma unroll
                for (int w = 1; w < W; w++) {
                    const int v = s_red[buf][w][0], r = s_red[buf][w][1];
                    if (v > vmax) { vmax = v; rmin = r; }   // warps are in rank order: ties keep the earlier
                }
            } else {
                const int v = lane < W ? s_red[buf][lane][0] : INT_MIN;
                const int r = lane < W ? s_red[buf][lane][1] : INT_MAX;
                vmax = __reduce_max_sync(0xffffffffu, v);
                rmin = __reduce_min_sync(0xffffffffu, v == vmax ? r : INT_MAX);
            }
        }
        if (vmax == 0) break;   // block-uniform: every point coincides with a sample, the winner is index 0 from here on
        cur = rank_to_index(rmin, L, SH);
        if (tid == 0) idxs[j] = cur + base;
    }
    for (int q = j + tid; q < m; q += W * 32) idxs[q] = base;
    // the reference leaves the final min-distances in temp (caller scratch): keep that contract
#pragma unroll
    for (int i = 0; i < PPL; i++) {
        const int k = rank_to_index(tid * PPL + i, L, SH);
        if (temp && (tid * PPL + i) < (bs << SH) && k < n) temp[k] = pt[i];
    }
    if (new_xyz) {   // optional fused gather of the sampled coordinates
        __syncthreads();
        for (int j = tid; j < m; j += W * 32) {
            const int k = idxs[j] - base;
            new_xyz[j * 3 + 0] = sxyz[k * 3 + 0];
            new_xyz[j * 3 + 1] = sxyz[k * 3 + 1];
            new_xyz[j * 3 + 2] = sxyz[k * 3 + 2];
        }
    }
}

// ---- one WARP per cloud (rank space <= 512): no block barrier, no shared-memory exchange ---------------------
// Each lane owns P consecutive ranks.  A round is: distance + running minimum per owned point (7 flops), a
// binary max tree over the lane's P minima (P-1 FMNMX, the tree is kept), one REDUX for the warp maximum, a
// first-match DESCENT of the kept tree (P-1-log2 P selects: the winner lane's first slot holding the maximum, i.e.
// the reference's tie order inside the lane) and one REDUX for the smallest winning rank.  ~10 instructions per
// point and round instead of ~20 for the four-warp geometry with per-element argmax tracking.
// Once the maximum is +0 every point coincides with a sampled one (pooled clouds wrap their points):
// the reference keeps returning index 0 from then on, so the remaining indices are 0 and the rounds stop.
template <int N>
__device__ __forceinline__ float pick_pow2(const float *v, int p) {   // v[p], p in [0, N), v in registers
    if constexpr (N == 1) {
        return v[0];
    } else {
        const float lo = pick_pow2<N / 2>(v, p), hi = pick_pow2<N / 2>(v + N / 2, p);
        return (p & (N / 2)) ? hi : lo;
    }
}

template <int P, int LEV>
__device__ __forceinline__ int descend(const float (&t)[2 * P], int vmax, int p) {
    if constexpr ((1 << LEV) > P) {
        return p;
    } else {
        constexpr int H = 1 << (LEV - 1);      // nodes of level LEV-1 = left children of level LEV
        float ev[H];
#pragma unroll
        for (int q = 0; q < H; q++) ev[q] = t[(1 << LEV) + 2 * q];
        const float left = pick_pow2<H>(ev, p);
        return descend<P, LEV + 1>(t, vmax, 2 * p + (__float_as_int(left) != vmax ? 1 : 0));
    }
}

constexpr int kFpsWarpsPerCta = 2;

template <int P>
__global__ void __launch_bounds__(kFpsWarpsPerCta * 32, 9)   // 18 warps / SM: the 2560 clouds of a 64-frame step are one wave
fps_warp_kernel(const float *__restrict__ xyz, int xs, float *__restrict__ temp, int *__restrict__ idxs,
                float *__restrict__ new_xyz, int b, int n, int m, int L, int SH, int f4_per_warp) {
    extern __shared__ float4 fps_sm4[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int cloud = blockIdx.x * kFpsWarpsPerCta + w;
    if (cloud >= b) return;                         // warps are independent: no block-level barrier below
    float4 *sp = fps_sm4 + (size_t)w * f4_per_warp;  // [n] staged cloud, then [m] sampled indices
    int *sidx = reinterpret_cast<int *>(sp + n);
    xyz += (size_t)cloud * n * xs; idxs += (size_t)cloud * m;
    if (temp) temp += (size_t)cloud * n;
    if (new_xyz) new_xyz += (size_t)cloud * m * 3;

    if (xs == 4 && (reinterpret_cast<uintptr_t>(xyz) & 15) == 0) {
        for (int k = lane; k < n; k += 32) sp[k] = __ldg(reinterpret_cast<const float4 *>(xyz) + k);
    } else {
        for (int k = lane; k < n; k += 32)
            sp[k] = make_float4(xyz[(size_t)k * xs], xyz[(size_t)k * xs + 1], xyz[(size_t)k * xs + 2], 0.f);
    }
    __syncwarp();

    const int bs = 1 << L;
    float px[P], py[P], pz[P], t[2 * P];            // t[P + i] = running minimum of rank lane*P + i; t[1..P-1] = max tree
#pragma unroll
    for (int i = 0; i < P; i++) {
        const int r = lane * P + i, k = rank_to_index(r, L, SH);
        const bool valid = r < (bs << SH) && k < n;
        const float4 q = sp[valid ? k : 0];
        px[i] = valid ? q.x : 0.f; py[i] = valid ? q.y : 0.f; pz[i] = valid ? q.z : 0.f;
        t[P + i] = valid ? (temp ? temp[k] : 1e10f) : -1.f;
    }
    int cur = 0, j = 1;
    if (lane == 0) sidx[0] = 0;
    for (; j < m; j++) {
        const float4 c = sp[cur];
#pragma unroll
        for (int i = 0; i < P; i++) {
            const float d = sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z));
            t[P + i] = fminf(d, t[P + i]);
        }
#pragma unroll
        for (int q = P - 1; q >= 1; q--) t[q] = fmaxf(t[2 * q], t[2 * q + 1]);
        // distances are >= 0 (or the -1 filler): their int bit patterns order like the floats
        const int vb = __float_as_int(P > 1 ? t[1] : t[P]);
        const int vmax = __reduce_max_sync(0xffffffffu, vb);
        if (vmax == 0) break;                       // every point already coincides with a sample: index 0 from here on
        const int slot = descend<P, 1>(t, vmax, 0);
        const int rmin = __reduce_min_sync(0xffffffffu, vb == vmax ? lane * P + slot : INT_MAX);
        cur = rank_to_index(rmin, L, SH);
        if (lane == 0) sidx[j] = cur;
    }
    for (int q = j + lane; q < m; q += 32) sidx[q] = 0;
    __syncwarp();
    for (int q = lane; q < m; q += 32) idxs[q] = sidx[q];
    if (new_xyz) {
        for (int e = lane; e < m * 3; e += 32) {
            const int q = e / 3, cc = e - q * 3;
            const float4 v = sp[sidx[q]];
            new_xyz[e] = cc == 0 ? v.x : (cc == 1 ? v.y : v.z);
        }
    }
    if (temp) {   // the reference leaves the final min-distances in temp (caller scratch): keep that contract
#pragma unroll
        for (int i = 0; i < P; i++) {
            const int r = lane * P + i, k = rank_to_index(r, L, SH);
            if (r < (bs << SH) && k < n) temp[k] = t[P + i];
        }
    }
}

// Any-size path: 1024 threads, cloud and temp stay in global/L2 (same traffic pattern as the
// reference) but the 10-barrier tree is replaced by REDUX + one barrier.  bs is 1024 for every
// n >= 1024 in both layouts, so thread t owns k = t, t+1024, ... exactly like the reference.
__global__ void __launch_bounds__(1024)
fps_global_kernel(const float *__restrict__ xyz, int xs, float *__restrict__ temp, int *__restrict__ idxs,
                  float *__restrict__ new_xyz, const int *__restrict__ xyz_batch_cnt,
                  const int *__restrict__ num_sampled, int n_uniform, int m_uniform, int L_uniform, int stack) {
    __shared__ int s_red[2][32][2];
    __shared__ int s_win[2];
    __shared__ long long s_off[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bi = blockIdx.x;
    int n, m, L;
    long long start, istart;
    if (stack) {
        if (tid == 0) {
            long long a = 0, c = 0;
            for (int k = 0; k < bi; k++) { a += xyz_batch_cnt[k]; c += num_sampled[k]; }
            s_off[0] = a; s_off[1] = c;
        }
        __syncthreads();
        start = s_off[0]; istart = s_off[1];
        n = xyz_batch_cnt[bi]; m = num_sampled[bi]; L = 10;
    } else {
        n = n_uniform; m = m_uniform; L = L_uniform;
        start = (long long)bi * n; istart = (long long)bi * m;
    }
    if (m <= 0) return;
    xyz += start * xs; temp += start; idxs += istart;
    if (new_xyz) new_xyz += istart * 3;
    const int base = stack ? (int)start : 0;
    const int bs = 1 << L;
    // rank of this thread among threads = bit-reversed id; threads >= bs own nothing
    const int trank = (tid < bs) ? (L == 0 ? 0 : (int)(__brev((unsigned)tid) >> (32 - L))) : INT_MAX;

    int cur = 0;
    if (tid == 0) idxs[0] = base;
    for (int j = 1; j < m; j++) {
        const float cx = xyz[(size_t)cur * xs + 0], cy = xyz[(size_t)cur * xs + 1], cz = xyz[(size_t)cur * xs + 2];
        float best = -1.f;
        int besti = 0;
        if (tid < bs)
            for (int k = tid; k < n; k += bs) {
                const float d = sqdist(__fsub_rn(xyz[(size_t)k * xs + 0], cx), __fsub_rn(xyz[(size_t)k * xs + 1], cy),
                                       __fsub_rn(xyz[(size_t)k * xs + 2], cz));
                const float d2 = fminf(d, temp[k]);
                temp[k] = d2;
                if (d2 > best) { best = d2; besti = k; }
            }
        const int vb = __float_as_int(best);
        int vmax = __reduce_max_sync(0xffffffffu, vb);
        int rmin = __reduce_min_sync(0xffffffffu, vb == vmax ? trank : INT_MAX);
        int kwin = __shfl_sync(0xffffffffu, besti, __ffs(__ballot_sync(0xffffffffu, vb == vmax && trank == rmin)) - 1);
        const int buf = j & 1;
        if (lane == 0) { s_red[buf][warp][0] = vmax; s_red[buf][warp][1] = rmin; }
        __syncthreads();
        const int v = s_red[buf][lane][0], r = s_red[buf][lane][1];
        vmax = __reduce_max_sync(0xffffffffu, v);
        rmin = __reduce_min_sync(0xffffffffu, v == vmax ? r : INT_MAX);
        const int wwin = __ffs(__ballot_sync(0xffffffffu, v == vmax && r == rmin)) - 1;
        // broadcast the winning warp's point index (double-buffered: the slot is rewritten two
        // barriers later; temp[k] is only ever touched by its owning thread)
        if (warp == wwin && lane == 0) s_win[buf] = kwin;
        __syncthreads();
        cur = s_win[buf];
        if (tid == 0) idxs[j] = cur + base;
    }
    if (new_xyz) {
        __syncthreads();
        for (int j = tid; j < m; j += 1024) {
            const int k = idxs[j] - base;
            new_xyz[j * 3 + 0] = xyz[(size_t)k * xs + 0];
            new_xyz[j * 3 + 1] = xyz[(size_t)k * xs + 1];
            new_xyz[j * 3 + 2] = xyz[(size_t)k * xs + 2];
        }
    }
}

// ---- FPS + ball query in one pass (power-of-two clouds of 32*P points, one warp per cloud) -------------------------
// Round j of the sampling measures every point of the cloud against sample j-1 - and those are exactly the distances
// the ball query around centroid j-1 tests against radius^2 (same operands up to sign, same contraction, hence the
// same bits).  So the query rides on the sampling round: one ballot per 32-point block on the distance the round
// has in a register anyway, a prefix of popcounts over the blocks in index order, an ordered append into a
// K-entry shared-memory row, one coalesced store of the row.  The centroids never travel through global memory
// and the second kernel's 7 flops per (centroid, point) pair disappear.
// Rank space = bit reversal (SH = 0, bs = n): rank brev5(l) * P + i is index brev_p(i) * 32 + l, so with the lanes
// permuted that way slot i is the 32-point block brev_p(i) in index order.
template <int PB>
__device__ __forceinline__ constexpr int brev_bits(int v) {
    int r = 0;
    for (int q = 0; q < PB; q++) r |= ((v >> q) & 1) << (PB - 1 - q);
    return r;
}
template <int P> struct Log2P { static constexpr int v = 1 + Log2P<P / 2>::v; };
template <> struct Log2P<1> { static constexpr int v = 0; };

__device__ __forceinline__ int brev_bits_rt(int v, int bits) { return bits == 0 ? 0 : (int)(__brev((unsigned)v) >> (32 - bits)); }

template <int P>
__global__ void __maxnreg__(112)   // 9 CTAs x 64 threads x 112 registers = 63 K: one wave of 2560 clouds
fps_bq_warp_kernel(const float *__restrict__ xyz, int xs, int *__restrict__ idxs, float *__restrict__ new_xyz,
                   int *__restrict__ ball_idx, float radius, int K, int b, int m, int f4_per_warp,
                   int *__restrict__ live_cnt, int dup_rows) {
    constexpr int n = 32 * P, L = 5 + Log2P<P>::v;
    extern __shared__ float4 fps_sm4[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int cloud = blockIdx.x * kFpsWarpsPerCta + w;
    if (cloud >= b) return;
    const int cloud_id = cloud;
    float4 *sp = fps_sm4 + (size_t)w * f4_per_warp;  // [n] staged cloud, [m] sampled indices, 2 x [K] ball rows
    int *sidx = reinterpret_cast<int *>(sp + n);
    int *rows = sidx + m;
    xyz += (size_t)cloud * n * xs; idxs += (size_t)cloud * m;
    new_xyz += (size_t)cloud * m * 3;
    ball_idx += (size_t)cloud * m * K;

    if (xs == 4 && (reinterpret_cast<uintptr_t>(xyz) & 15) == 0) {
        for (int k = lane; k < n; k += 32) sp[k] = __ldg(reinterpret_cast<const float4 *>(xyz) + k);
    } else {
        for (int k = lane; k < n; k += 32)
            sp[k] = make_float4(xyz[(size_t)k * xs], xyz[(size_t)k * xs + 1], xyz[(size_t)k * xs + 2], 0.f);
    }
    __syncwarp();

    // lane l owns the ranks brev5(l) * P + i, i.e. the indices brev_p(i) * 32 + l: inside a 32-point block the lanes
    // sit in index order, so a ballot over slot i IS the index-ordered hit mask of block brev_p(i)
    const int rank0 = (int)(__brev((unsigned)lane) >> 27) * P;
    const float r2 = __fmul_rn(radius, radius);
    float px[P], py[P], pz[P], t[2 * P];
#pragma unroll
    for (int i = 0; i < P; i++) {
        const float4 q = sp[rank_to_index(rank0 + i, L, 0)];
        px[i] = q.x; py[i] = q.y; pz[i] = q.z;
        t[P + i] = 1e10f;
    }
    constexpr int PB = Log2P<P>::v;
    const int my_slot = lane < P ? brev_bits_rt(lane, PB) : -1;     // the register slot that is block `lane`
    int buf = 0;
    // The ball of one centroid from the P hit masks: lane s < P takes block s, an inclusive scan of the popcounts
    // gives its first slot, then it appends its own hits (index order = bit order) until the row is full.  The row
    // leaves as one coalesced store padded with its first entry; empty rows are (-1, 0, 0, ...) as fill = 1 of
    // mgar_ball_query_stack.  Rows are double-buffered: the barrier of round j+1 orders round j's reads before
    // round j+2's writes.
    auto emit = [&](int c, unsigned mask) {
        int *row = rows + buf * K;
        int incl = __popc(mask);
#pragma unroll
        for (int o = 1; o < P; o *= 2) {
            const int up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        const int cnt = __shfl_sync(0xffffffffu, incl, P - 1);
        int pos = incl - __popc(mask);
        while (mask && pos < K) {
            row[pos++] = lane * 32 + __ffs(mask) - 1;
            mask &= mask - 1;
        }
        __syncwarp();
        const int first = row[0];
        const int v0 = lane < K ? row[lane] : 0, v1 = lane + 32 < K ? row[lane + 32] : 0;
        int *out = ball_idx + (size_t)c * K;
        if (lane < K) out[lane] = cnt == 0 ? (lane == 0 ? -1 : 0) : (lane < cnt ? v0 : first);
        if (lane + 32 < K) out[lane + 32] = cnt == 0 ? 0 : (lane + 32 < cnt ? v1 : first);
        buf ^= 1;
    };
    int cur = 0, j = 1;
    if (lane == 0) sidx[0] = 0;
    bool done = false;                               // every point coincides with a sample: index 0 from here on
    for (; j < m; j++) {
        const float4 c = sp[cur];
        unsigned mask = 0;                           // lane s < P keeps the hit mask of block s, the others nothing
#pragma unroll
        for (int i = 0; i < P; i++) {
            const float d = sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z));
            const unsigned hb = __ballot_sync(0xffffffffu, d < r2);
            mask = my_slot == i ? hb : mask;
            t[P + i] = fminf(d, t[P + i]);
        }
#pragma unroll
        for (int q = P - 1; q >= 1; q--) t[q] = fmaxf(t[2 * q], t[2 * q + 1]);
        const int vb = __float_as_int(P > 1 ? t[1] : t[P]);
        const int vmax = __reduce_max_sync(0xffffffffu, vb);
        const int slot = descend<P, 1>(t, vmax, 0);
        const int rmin = __reduce_min_sync(0xffffffffu, vb == vmax ? rank0 + slot : INT_MAX);
        emit(j - 1, mask);                           // the ball of centroid j-1, while the REDUX is in flight
        if (vmax == 0) { done = true; break; }
        cur = rank_to_index(rmin, L, 0);
        if (lane == 0) sidx[j] = cur;
    }
    if (!done) {                                     // the last sample has had no round of its own
        const float4 c = sp[cur];
        unsigned mask = 0;
#pragma unroll
        for (int i = 0; i < P; i++) {
            const unsigned hb = __ballot_sync(0xffffffffu, sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z)) < r2);
            mask = my_slot == i ? hb : mask;
        }
        emit(m - 1, mask);
    } else {                                         // samples j .. m-1 are point 0 = sample 0: their balls are row 0
        for (int q = j + lane; q < m; q += 32) sidx[q] = 0;
        int *row = rows;
        __syncwarp();
        for (int l = lane; l < K; l += 32) row[l] = ball_idx[l];   // written by this warp (same lanes) in round 1
        __syncwarp();
        // dup_rows == 0: the caller only walks the distinct centroids (live_cnt), rounded up to eight
        // (a tile of the MLP kernel = 128 positions = 4 centroids of 32 samples, 8 of 16)
        const int lim = dup_rows ? m : min(m, (j + 7) & ~7);
        for (int e = lane; e < (lim - j) * K; e += 32) ball_idx[(size_t)j * K + e] = row[e % K];
    }
    // number of DISTINCT samples: centroids live .. m-1 repeat centroid 0 (same ball, same features)
    if (live_cnt && lane == 0) live_cnt[cloud_id] = done ? j : m;
    __syncwarp();
    for (int q = lane; q < m; q += 32) idxs[q] = sidx[q];
    for (int e = lane; e < m * 3; e += 32) {
        const int q = e / 3, cc = e - q * 3;
        const float4 v = sp[sidx[q]];
        new_xyz[e] = cc == 0 ? v.x : (cc == 1 ? v.y : v.z);
    }
}

template <int W, int PPL>
static int launch_reg(int grid, size_t smem, cudaStream_t st, const float *xyz, int xs, float *temp, int *idx,
                      float *new_xyz, const int *cnt, const int *ms, int n, int m, int L, int stack) {
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(fps_reg_kernel<W, PPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { set_error("fps: smem opt-in failed: %s", cudaGetErrorString(e)); return MGAR_ECUDA; }
    }
    fps_reg_kernel<W, PPL><<<grid, W * 32, smem, st>>>(xyz, xs, temp, idx, new_xyz, cnt, ms, n, m, L, stack, (int)(smem / 12));
    count_launch();
    return check_launch("fps");
}

static int SH_of(int V, int L) { int sh = 0; while (((1 << L) << sh) < V) sh++; return sh; }
static int floor_log2(int v) { int l = 0; while ((2 << l) <= v) l++; return l; }

// V = padded rank space.  Geometry table: (warps, ranks per thread) chosen so one round is short
// for a single cloud yet several clouds still share an SM; overridable for tuning.
static int fps_dispatch(int grid, int V, int max_n, cudaStream_t st, const float *xyz, int xs, float *temp, int *idx,
                        float *new_xyz, const int *cnt, const int *ms, int n, int m, int L, int stack) {
    // many small clouds: one warp per cloud (throughput); a few clouds: several warps per cloud (latency)
    if (!stack && V <= 512 && grid >= 4 * kNumSMs && !getenv("MGAR_FPS_GEOM")) {
        int P = 1;
        while (32 * P < V) P *= 2;
        const int f4 = n + (m + 3) / 4;
        const size_t sm = (size_t)kFpsWarpsPerCta * f4 * sizeof(float4);
        if (sm <= 200 * 1024) {
            const int ctas = (grid + kFpsWarpsPerCta - 1) / kFpsWarpsPerCta;
#define MGAR_FPS_WARP(p)                                                                                              \
    if (P == p) {                                                                                                     \
        if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(fps_warp_kernel<p>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
        fps_warp_kernel<p><<<ctas, kFpsWarpsPerCta * 32, sm, st>>>(xyz, xs, temp, idx, new_xyz, grid, n, m, L, SH_of(V, L), f4); \
        count_launch();                                                                                               \
        return check_launch("fps(warp)");                                                                             \
    }
            MGAR_FPS_WARP(1) MGAR_FPS_WARP(2) MGAR_FPS_WARP(4) MGAR_FPS_WARP(8) MGAR_FPS_WARP(16)
#undef MGAR_FPS_WARP
        }
    }
    const size_t smem = (size_t)max_n * 3 * sizeof(float);
    int W, PPL;
    if (V <= 32) { W = 1; PPL = 1; }
    else if (V <= 64) { W = 1; PPL = 2; }
    else if (V <= 128) { W = 1; PPL = 4; }
    else if (V <= 256) { W = 2; PPL = 4; }
    else if (V <= 512) { W = 4; PPL = 4; }
    else if (V <= 1024) { W = 4; PPL = 8; }
    else if (V <= 2048) { W = 8; PPL = 8; }
    else if (V <= 4096) { W = 16; PPL = 8; }
    else { W = 32; PPL = 8; }
    if (const char *e = getenv("MGAR_FPS_GEOM")) {  // "W,PPL" tuning override (must cover V)
        int w = 0, p = 0;
        if (sscanf(e, "%d,%d", &w, &p) == 2 && w * 32 * p >= V) { W = w; PPL = p; }
    }
    if (V > 8192 || smem > 200 * 1024) {
        MGAR_REQUIRE(temp, "fps: clouds this large need the caller's temp scratch");
        fps_global_kernel<<<grid, 1024, 0, st>>>(xyz, xs, temp, idx, new_xyz, cnt, ms, n, m, L, stack);
        count_launch();
        return check_launch("fps(global)");
    }
#define MGAR_FPS_CASE(w, p) \
    if (W == w && PPL == p) return launch_reg<w, p>(grid, smem, st, xyz, xs, temp, idx, new_xyz, cnt, ms, n, m, L, stack);
    MGAR_FPS_CASE(1, 1) MGAR_FPS_CASE(1, 2) MGAR_FPS_CASE(1, 4) MGAR_FPS_CASE(1, 8) MGAR_FPS_CASE(1, 16)
    MGAR_FPS_CASE(2, 2) MGAR_FPS_CASE(2, 4) MGAR_FPS_CASE(2, 8) MGAR_FPS_CASE(2, 16)
    MGAR_FPS_CASE(4, 2) MGAR_FPS_CASE(4, 4) MGAR_FPS_CASE(4, 8) MGAR_FPS_CASE(4, 16)
    MGAR_FPS_CASE(8, 1) MGAR_FPS_CASE(8, 2) MGAR_FPS_CASE(8, 4) MGAR_FPS_CASE(8, 8) MGAR_FPS_CASE(8, 16)
    MGAR_FPS_CASE(16, 1) MGAR_FPS_CASE(16, 2) MGAR_FPS_CASE(16, 4) MGAR_FPS_CASE(16, 8)
    MGAR_FPS_CASE(32, 1) MGAR_FPS_CASE(32, 2) MGAR_FPS_CASE(32, 4) MGAR_FPS_CASE(32, 8)
#undef MGAR_FPS_CASE
    set_error("fps: unsupported geometry W=%d PPL=%d", W, PPL);
    return MGAR_EINVAL;
}

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_fps_batch(int b, int n, int m, const float *xyz, int xyz_stride, float *temp, int *idx,
                              float *new_xyz, void *stream) {
    MGAR_REQUIRE(b >= 0 && n >= 0 && m >= 0, "fps_batch: negative size");
    if (b == 0 || m == 0) return MGAR_OK;
    MGAR_REQUIRE(n >= 1, "fps_batch: n must be >= 1 when m > 0");
    MGAR_REQUIRE(xyz && idx, "fps_batch: null pointer");
    MGAR_REQUIRE(xyz_stride >= 3, "fps_batch: xyz_stride must be >= 3");
    int L = floor_log2(n);
    if (L > 10) L = 10;
    const int bs = 1 << L;
    int stripes = (n + bs - 1) / bs, SH = 0;
    while ((1 << SH) < stripes) SH++;
    return fps_dispatch(b, bs << SH, n, (cudaStream_t)stream, xyz, xyz_stride, temp, idx, new_xyz, nullptr, nullptr, n, m, L, 0);
}

extern "C" int mgar_fps_stack(int batch, int max_n, const float *xyz, float *temp, const int *xyz_batch_cnt,
                              int *idx, const int *num_sampled, void *stream) {
    MGAR_REQUIRE(batch >= 0 && max_n >= 0, "fps_stack: negative size");
    if (batch == 0) return MGAR_OK;
    MGAR_REQUIRE(xyz && temp && idx && xyz_batch_cnt && num_sampled, "fps_stack: null pointer");
    int stripes = (max_n + 1023) / 1024, SH = 0;
    if (stripes < 1) stripes = 1;
    while ((1 << SH) < stripes) SH++;
    return fps_dispatch(batch, 1024 << SH, max_n, (cudaStream_t)stream, xyz, 3, temp, idx, nullptr, xyz_batch_cnt,
                        num_sampled, 0, 0, 10, 1);
}

// ---- fused sampling + ball query: entry points --------------------------------------------------------------
// n in {128, 256, 512}: one CTA of W warps per cloud (fps_bq_cta.cuh), any number of clouds;
// n in {32, 64}: one warp per cloud (fps_bq_warp_kernel), >= 592 clouds.  MGAR_FPS_BQ=warp forces the warp kernel
// (A/B runs), MGAR_FPS_BQ_GEOM="W,PPL" picks another CTA geometry with 32 * W * PPL == n.
static bool bq_use_cta(int n) {
    if (n != 128 && n != 256 && n != 512) return false;
    const char *e = getenv("MGAR_FPS_BQ");
    return !(e && e[0] == 'w');
}

extern "C" int mgar_fps_ball_query_supported(int b, int n, int m, int nsample) {
    if (b < 1 || m < 1 || nsample < 1 || nsample > 64 || getenv("MGAR_NO_FPS_BQ")) return 0;
    if (bq_use_cta(n)) return (size_t)n * 16 + (size_t)m * 4 + 8 * 2 * (size_t)nsample * 4 <= 200 * 1024;
    if (b < 4 * kNumSMs) return 0;
    return n == 32 || n == 64 || n == 128 || n == 256 || n == 512;
}

extern "C" size_t mgar_fps_ball_query_workspace_bytes(int b) { return b > 0 ? ((size_t)b + 4) * sizeof(int) : 0; }   // order + n_heavy

// The library's side stream of the current device (the split launch below runs two kernels side by side): created on first
// use, one per device, never destroyed.  Not memory: the "no hidden allocation" rule of the boundary is about buffers.
namespace {
struct SideStream { cudaStream_t stream; cudaEvent_t fork, join; bool ok; };
SideStream *side_stream() {
    static SideStream table[64];
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("fps_ball_query: no current device"); return nullptr; }
    std::lock_guard<std::mutex> lock(mu);
    SideStream &s = table[dev];
    if (!s.ok) {
        if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming) != cudaSuccess) {
            set_error("fps_ball_query: could not create the side stream: %s", cudaGetErrorString(cudaGetLastError()));
            return nullptr;
        }
        s.ok = true;
    }
    return &s;
}
}  // namespace

template <int W, int PPL>
static int launch_bq_cta(int b, int m, const float *xyz, int xs, float radius, int K, int *fps_idx, float *new_xyz,
                         int *ball_idx, const int *order, int *live_cnt, int dup_rows, cudaStream_t st) {
    size_t sm = (size_t)(32 * W * PPL) * sizeof(float4) + (size_t)m * sizeof(int) + (size_t)W * 2 * K * sizeof(int);
    // Resident CTAs per SM are CAPPED (by asking for more shared memory than the CTA uses): every resident warp gets an
    // equal share of the issue slots, so a cloud's round lasts (resident warps per scheduler) x (its instructions) and the
    // step lasts at least as long as the longest cloud's 512 rounds; too few warps leave the schedulers idle instead.
    // Measured on the bench step (profiles/r02_fps_bq_sweep.txt): W=2: 4 / 6 / 8 / 10 / 14 CTAs per SM -> 0.55 / 0.38 /
    // 0.311 / 0.311 / 0.313 ms; W=4: 2 / 3 / 4 / 6 / 10 -> 0.95 / 0.54 / 0.41 / 0.33 / 0.314 ms.
    int ctas_per_sm = 20 / W;                                  // 5 warps per scheduler
    if (const char *e = getenv("MGAR_FPS_BQ_CTAS")) { const int v = atoi(e); if (v >= 1 && v <= 32) ctas_per_sm = v; }
    const size_t cap = ((size_t)227 * 1024) / ctas_per_sm - 1024;   // 1 KB per CTA is reserved by the driver
    if (b > kNumSMs * ctas_per_sm && cap > sm) sm = cap & ~(size_t)127;
    if (sm > 48 * 1024)
        MGAR_CUDA(cudaFuncSetAttribute(fps_bq_cta_kernel<W, PPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    fps_bq_cta_kernel<W, PPL><<<b, W * 32, sm, st>>>(xyz, xs, fps_idx, new_xyz, ball_idx, radius, K, b, m, order, live_cnt, dup_rows);
    count_launch();
    return check_launch("fps_ball_query(cta)");
}

extern "C" int mgar_fps_ball_query(int b, int n, int m, const float *xyz, int xyz_stride, float radius, int nsample,
                                   int *fps_idx, float *new_xyz, int *ball_idx, int *live_cnt, int write_dup_rows,
                                   const int *cloud_weight, void *workspace, size_t workspace_bytes, void *stream) {
    MGAR_REQUIRE(mgar_fps_ball_query_supported(b, n, m, nsample), "fps_ball_query: unsupported shape b=%d n=%d m=%d nsample=%d "
                 "(ask mgar_fps_ball_query_supported first)", b, n, m, nsample);
    MGAR_REQUIRE(xyz && fps_idx && new_xyz && ball_idx, "fps_ball_query: null pointer");
    MGAR_REQUIRE(xyz_stride >= 3, "fps_ball_query: xyz_stride must be >= 3");
    cudaStream_t st = (cudaStream_t)stream;
    if (bq_use_cta(n)) {
        const int *order = nullptr;
        bool order_pending = false;                  // the sort is launched by whichever path below runs
        if (cloud_weight && b > kNumSMs) {          // longest clouds first (with <= one CTA per SM the order is moot)
            if (!workspace || workspace_bytes < mgar_fps_ball_query_workspace_bytes(b)) {
                set_error("fps_ball_query: cloud_weight needs a workspace of mgar_fps_ball_query_workspace_bytes(b) = %zu bytes",
                          mgar_fps_ball_query_workspace_bytes(b));
                return MGAR_ENOMEM;
            }
            order = static_cast<const int *>(workspace);
            order_pending = true;
        }
        // pooled clouds are periodic (slot k = point k mod count): with the count as a (verified) hint only the distinct
        // blocks take part in a round (fps_bq_dup.cuh).  Bit-identical, 35 % fewer point-rounds - and on its own SLOWER on the
        // bench step (0.338 ms at its best residency against 0.309 ms, profiles/r02_fps_bq_sweep.txt): the step lasts as long
        // as the 512-round clouds, whose round is latency-bound and got a few instructions longer.  MGAR_FPS_BQ_DUP=1 runs
        // it alone (the record of that experiment).
        auto sort_clouds = [&](int *n_heavy, int heavy_from) -> int {
            order_clouds_kernel<<<1, 1024, 0, st>>>(cloud_weight, b, static_cast<int *>(workspace), n_heavy, heavy_from);
            count_launch();
            return check_launch("fps_ball_query(order)");
        };
        // (radius < 1e9: the periodic kernel parks its unused register slots 1e18 away, beyond any sane radius)
        const bool dup_ok = n == kDupN && cloud_weight && order && !getenv("MGAR_FPS_BQ_GEOM") && radius < 1e9f;
        auto dup_smem = [&](int ctas_per_sm) {
            size_t sm = (size_t)kDupN * sizeof(float4) + (size_t)m * sizeof(int) + (size_t)kDupW * 2 * nsample * sizeof(int);
            if (const char *e = getenv("MGAR_FPS_BQ_CTAS")) { const int v = atoi(e); if (v >= 1 && v <= 32) ctas_per_sm = v; }
            const size_t cap = ((size_t)227 * 1024) / ctas_per_sm - 1024;
            if (b > kNumSMs * ctas_per_sm && cap > sm) sm = cap & ~(size_t)127;
            return sm;
        };
        if (dup_ok && getenv("MGAR_FPS_BQ_DUP")) {
            if (int rc = sort_clouds(nullptr, 0)) return rc;
            const size_t sm = dup_smem(10);
            if (sm > 48 * 1024)
                MGAR_CUDA(cudaFuncSetAttribute(fps_bq_dup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
            fps_bq_dup_kernel<<<b, kDupW * 32, sm, st>>>(xyz, xyz_stride, fps_idx, new_xyz, ball_idx, radius, nsample, b, m, order,
                                                         cloud_weight, live_cnt, write_dup_rows);
            count_launch();
            return check_launch("fps_ball_query(dup)");
        }
        // SPLIT (the default for many wrapped 512-slot clouds): the two kernels side by side, each on the clouds it is good
        // at.  The clouds with >= split_from distinct points (the first n_heavy entries of `order`) go to the four-warp
        // kernel, whose round is the shortest (511 against 771 cycles with the SM to itself,
        // profiles/r02_fps_round_latency.json) - they are the step's critical path; all the others go to the periodic-cloud
        // kernel, which spends instructions only on their distinct blocks - they are the step's instruction count.  The
        // second launch runs on a side stream of the library (fork / join by events, legal under stream capture too).
        int split_from = 320;                    // sweeps: profiles/r02_fps_split_sweep.txt
        if (const char *e = getenv("MGAR_FPS_BQ_SPLIT")) split_from = atoi(e);
        if (dup_ok && split_from > 0 && b >= 8 * kNumSMs) {
            SideStream *side = side_stream();
            if (!side) return MGAR_ECUDA;
            int *n_heavy = static_cast<int *>(workspace) + b;
            if (int rc = sort_clouds(n_heavy, split_from)) return rc;
            MGAR_CUDA(cudaEventRecord(side->fork, st));
            MGAR_CUDA(cudaStreamWaitEvent(side->stream, side->fork, 0));
            {   // long clouds: four warps each, on the caller's stream (MGAR_FPS_BQ_LONG=2: two warps, for A/B runs)
                const char *lg = getenv("MGAR_FPS_BQ_LONG");
                const int lw = (lg && lg[0] == '2') ? 2 : 4;
                const size_t sm = (size_t)kDupN * sizeof(float4) + (size_t)m * sizeof(int) + (size_t)lw * 2 * nsample * sizeof(int);
                if (lw == 4) {
                    if (sm > 48 * 1024)
                        MGAR_CUDA(cudaFuncSetAttribute(fps_bq_cta_kernel<4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
                    // both kernels ask for the same shared-memory carve-out: CTAs of kernels with different L1 / shared splits
                    // do not share an SM, and the two launches would run one after the other
                    MGAR_CUDA(cudaFuncSetAttribute(fps_bq_cta_kernel<4, 4>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                                   (int)cudaSharedmemCarveoutMaxShared));
                    if (getenv("MGAR_FPS_BQ_NO_REDUX")) {
                        fps_bq_cta_kernel<4, 4><<<b, 128, sm, st>>>(xyz, xyz_stride, fps_idx, new_xyz, ball_idx, radius, nsample, b, m, order,
                                                                    live_cnt, write_dup_rows, n_heavy);
                    } else {
                        MGAR_CUDA(cudaFuncSetAttribute(fps_bq_cta_kernel<4, 4, true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                                       (int)cudaSharedmemCarveoutMaxShared));
                        if (sm > 48 * 1024)
                            MGAR_CUDA(cudaFuncSetAttribute(fps_bq_cta_kernel<4, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
                        fps_bq_cta_kernel<4, 4, true><<<b, 128, sm, st>>>(xyz, xyz_stride, fps_idx, new_xyz, ball_idx, radius, nsample, b, m,
                                                                          order, live_cnt, write_dup_rows, n_heavy);
                    }
                } else {
                    if (sm > 48 * 1024)
                        MGAR_CUDA(cudaFuncSetAttribute(fps_bq_cta_kernel<2, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
                    MGAR_CUDA(cudaFuncSetAttribute(fps_bq_cta_kernel<2, 8>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                                   (int)cudaSharedmemCarveoutMaxShared));
                    fps_bq_cta_kernel<2, 8><<<b, 64, sm, st>>>(xyz, xyz_stride, fps_idx, new_xyz, ball_idx, radius, nsample, b, m, order,
                                                               live_cnt, write_dup_rows, n_heavy);
                }
                count_launch();
                if (int rc = check_launch("fps_ball_query(long clouds)")) return rc;
            }
            {   // every other cloud: the periodic-cloud kernel, on the side stream
                const size_t sm = dup_smem(10);
                if (sm > 48 * 1024)
                    MGAR_CUDA(cudaFuncSetAttribute(fps_bq_dup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
                MGAR_CUDA(cudaFuncSetAttribute(fps_bq_dup_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                               (int)cudaSharedmemCarveoutMaxShared));
                fps_bq_dup_kernel<<<b, kDupW * 32, sm, side->stream>>>(xyz, xyz_stride, fps_idx, new_xyz, ball_idx, radius, nsample, b,
                                                                       m, order, cloud_weight, live_cnt, write_dup_rows, n_heavy);
                count_launch();
                if (int rc = check_launch("fps_ball_query(wrapped clouds)")) return rc;
            }
            MGAR_CUDA(cudaEventRecord(side->join, side->stream));
            MGAR_CUDA(cudaStreamWaitEvent(st, side->join, 0));
            return MGAR_OK;
        }
        // Two warps per cloud cost the fewest instructions (throughput: many clouds per SM); four warps have the shorter
        // round (latency: 511 against 771 cycles per round when a cloud has the SM to itself,
        // profiles/r02_fps_round_latency.json) - used while there are fewer than 8 clouds per SM (a single frame: 40).
        if (order_pending) if (int rc = sort_clouds(nullptr, 0)) return rc;
        int W = b < 8 * kNumSMs ? 4 : 2, PPL = n / (32 * W);
        if (const char *e = getenv("MGAR_FPS_BQ_GEOM")) {
            int w = 0, p = 0;
            if (sscanf(e, "%d,%d", &w, &p) == 2 && 32 * w * p == n) { W = w; PPL = p; }
        }
#define MGAR_BQ_CTA(w, p)                                                                                               \
    if (W == w && PPL == p)                                                                                             \
        return launch_bq_cta<w, p>(b, m, xyz, xyz_stride, radius, nsample, fps_idx, new_xyz, ball_idx, order, live_cnt, \
                                   write_dup_rows, st);
        MGAR_BQ_CTA(4, 4) MGAR_BQ_CTA(2, 8) MGAR_BQ_CTA(8, 2) MGAR_BQ_CTA(4, 2) MGAR_BQ_CTA(2, 4) MGAR_BQ_CTA(8, 1)
        MGAR_BQ_CTA(4, 1) MGAR_BQ_CTA(2, 2) MGAR_BQ_CTA(1, 4)
#undef MGAR_BQ_CTA
        set_error("fps_ball_query: no kernel for W=%d PPL=%d", W, PPL);
        return MGAR_EINVAL;
    }
    const int f4 = n + (m + 2 * nsample + 3) / 4;
    const size_t sm = (size_t)kFpsWarpsPerCta * f4 * sizeof(float4);
    MGAR_REQUIRE(sm <= 200 * 1024, "fps_ball_query: m=%d too large for shared memory", m);
    const int ctas = (b + kFpsWarpsPerCta - 1) / kFpsWarpsPerCta;
#define MGAR_FPS_BQ(p)                                                                                               \
    if (n == 32 * p) {                                                                                               \
        if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(fps_bq_warp_kernel<p>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
        fps_bq_warp_kernel<p><<<ctas, kFpsWarpsPerCta * 32, sm, st>>>(xyz, xyz_stride, fps_idx, new_xyz, ball_idx, radius, nsample, b, m, f4, \
                                                                      live_cnt, write_dup_rows);                     \
        count_launch();                                                                                              \
        return check_launch("fps_ball_query");                                                                       \
    }
    MGAR_FPS_BQ(1) MGAR_FPS_BQ(2) MGAR_FPS_BQ(4) MGAR_FPS_BQ(8) MGAR_FPS_BQ(16)
#undef MGAR_FPS_BQ
    return MGAR_EINVAL;
}
