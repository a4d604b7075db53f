// FPS + ball query in one pass, ONE CTA OF W WARPS PER CLOUD (n = 32 * W * PPL points, n <= 512).
//
// Same contract and the same bits as fps_bq_warp_kernel (fps.cu): round j of the sampling measures every point
// against sample j-1, and those distances are the ball-query tests around centroid j-1 (farthest_point_sampling_kernel
// sampling_gpu.cu:25-140 + ball_query_kernel_stack ball_query_gpu.cu:16-66).  What changes is the shape of the work:
//
//   * the warp-per-cloud kernel is a chain of ~300 dependent-ish instructions per round in ONE warp (~2000 cycles per
//     round whatever else the SM does), and a step lasts as long as its longest cloud: 512 rounds x 2000 cycles for the
//     7 % of the persons that pooled >= 512 points, while the median person (145 distinct points) is done after 145
//     rounds.  Here a cloud's round is spread over W warps (PPL points per lane): the chain is one distance + a
//     (PPL-1)-deep max tree + two REDUX + one 8-byte exchange through shared memory and ONE block barrier.
//   * the ball rows are not emitted inside the round: a round only stores its 32-point hit masks (one STS per warp)
//     into a shared-memory ring; every kRing rounds the W warps turn the ring into rows - two centroids per warp pass
//     (half-warp per centroid: 16 lanes = the 16 blocks of a cloud), so the serial append loop and the row store are
//     off the round's critical path and amortised.
//   * clouds are taken in the caller's `order` (longest first: the pooled point count is known before the sampling),
//     so the short clouds fill in behind the long ones instead of the long ones defining a tail.
//
// Tie order (see the header of fps.cu): the winner among equal maxima minimises rank(k) = brev_L(k), L = log2 n.
// Lane l of warp w holds, in register slot i, the point k = blk * 32 + l of block blk = brev_NB(w * PPL + i)
// (NB = log2(W * PPL) block bits): then rank(k) = brev5(l) << NB | (w * PPL + i) - increasing in the slot inside a lane,
// so "first slot holding the maximum" is the lane's winner, one REDUX.MIN over the ranks the warp's, and the smallest
// rank over the W exchanged (value, rank) pairs the cloud's.  A ballot over slot i is the hit mask of block blk in index
// order; the ring stores it at position w * PPL + i and the emitter reads position brev_NB(s) for block s.
#pragma once
#include <limits.h>

#include "mgar_common.cuh"

namespace mgar {

constexpr int kBqRing = 8;      // rounds per emission batch (2 centroids per warp pass: W = 4 -> one pass per batch)

template <int BITS>
__device__ __forceinline__ int brev_n(int v) { return BITS == 0 ? 0 : (int)(__brev((unsigned)v) >> (32 - BITS)); }

template <int V> struct CtLog2 { static constexpr int v = 1 + CtLog2<V / 2>::v; };
template <> struct CtLog2<1> { static constexpr int v = 0; };

template <int W, int PPL>
__global__ void __launch_bounds__(W * 32)
fps_bq_cta_kernel(const float *__restrict__ xyz, int xs, int *__restrict__ idxs, float *__restrict__ new_xyz,
                  int *__restrict__ ball_idx, float radius, int K, int b, int m, const int *__restrict__ order,
                  int *__restrict__ live_cnt, int dup_rows) {
    constexpr int NBLK = W * PPL, NB = CtLog2<NBLK>::v, n = 32 * NBLK, L = 5 + NB;
    static_assert(NBLK <= 16 && (NBLK & (NBLK - 1)) == 0, "a cloud is at most 16 blocks of 32 points");
    extern __shared__ float4 bq_sm4[];
    __shared__ unsigned long long s_key[2][W];
    __shared__ unsigned s_ring[2][kBqRing][16];
    __shared__ int s_row0[64];
    float4 *sp = bq_sm4;                                   // [n] staged cloud
    int *sidx = reinterpret_cast<int *>(sp + n);           // [m] sampled indices
    int *rows = sidx + m;                                  // [W][2][K] rows being assembled

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int cloud = order ? __ldg(order + blockIdx.x) : (int)blockIdx.x;
    if (cloud < 0 || cloud >= b) return;
    xyz += (size_t)cloud * n * xs; idxs += (size_t)cloud * m;
    new_xyz += (size_t)cloud * m * 3;
    ball_idx += (size_t)cloud * m * K;

    if (xs == 4 && (reinterpret_cast<uintptr_t>(xyz) & 15) == 0) {
        for (int k = tid; k < n; k += W * 32) sp[k] = __ldg(reinterpret_cast<const float4 *>(xyz) + k);
    } else {
        for (int k = tid; k < n; k += W * 32)
            sp[k] = make_float4(xyz[(size_t)k * xs], xyz[(size_t)k * xs + 1], xyz[(size_t)k * xs + 2], 0.f);
    }
    __syncthreads();

    const float r2 = __fmul_rn(radius, radius);
    const int rank0 = (brev_n<5>(lane) << NB) | (w * PPL);
    float px[PPL], py[PPL], pz[PPL], t[PPL];
#pragma unroll
    for (int i = 0; i < PPL; i++) {
        const float4 q = sp[brev_n<NB>(w * PPL + i) * 32 + lane];
        px[i] = q.x; py[i] = q.y; pz[i] = q.z;
        t[i] = 1e10f;
    }

    // ---- rows of the centroids [c_emitted, c_end) from the ring (same batch: c_end - c_emitted <= kBqRing) ----------
    int c_emitted = 0;
    auto emit_batch = [&](int c_end) {
        const int c0 = c_emitted, buf = (c0 / kBqRing) & 1;
        const int half = lane >> 4, s = lane & 15;
        int *row = rows + (w * 2 + half) * K;
        for (int rr = 2 * w; rr < kBqRing && c0 + rr < c_end; rr += 2 * W) {
            const int c = c0 + rr + half;
            const bool valid = c < c_end;
            unsigned mask = (valid && s < NBLK) ? s_ring[buf][rr + half][brev_n<NB>(s)] : 0u;
            const int pc = __popc(mask);
            int incl = pc;
#pragma unroll
            for (int o = 1; o < 16; o *= 2) {
                const int up = __shfl_up_sync(0xffffffffu, incl, o, 16);
                if (s >= o) incl += up;
            }
            int pos = incl - pc;
            while (mask && pos < K) {                      // this block's hits, in index (= bit) order
                row[pos++] = s * 32 + __ffs(mask) - 1;
                mask &= mask - 1;
            }
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; h++) {                  // the two rows leave as coalesced stores of the whole warp
                const int ch = c0 + rr + h;
                if (ch >= c_end) break;
                const int cnt = __shfl_sync(0xffffffffu, incl, h * 16 + 15);
                const int *rh = rows + (w * 2 + h) * K;
                const int first = rh[0];
                for (int l = lane; l < K; l += 32) {
                    const int v = cnt == 0 ? (l == 0 ? -1 : 0) : (l < cnt ? rh[l] : first);
                    ball_idx[(size_t)ch * K + l] = v;
                    if (ch == 0) s_row0[l] = v;
                }
            }
            __syncwarp();
        }
        c_emitted = c_end;
    };

    int cur = 0, j = 1;
    bool done = false;                                    // every point coincides with a sample: index 0 from here on
    if (tid == 0) sidx[0] = 0;
    for (; j < m; j++) {
        const float4 c = sp[cur];
        unsigned hb[PPL];
#pragma unroll
        for (int i = 0; i < PPL; i++) {
            const float d = sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z));
            hb[i] = __ballot_sync(0xffffffffu, d < r2);
            t[i] = fminf(d, t[i]);
        }
        float mx = t[0];
#pragma unroll
        for (int i = 1; i < PPL; i++) mx = fmaxf(mx, t[i]);
        int slot = PPL - 1;
#pragma unroll
        for (int i = PPL - 2; i >= 0; i--) slot = t[i] == mx ? i : slot;
        // distances are >= 0: their int bit patterns order like the floats
        const int vb = __float_as_int(mx);
        const int vmax_w = __reduce_max_sync(0xffffffffu, vb);
        const int rmin_w = __reduce_min_sync(0xffffffffu, vb == vmax_w ? rank0 + slot : INT_MAX);
        if (lane == 0) {
            s_key[j & 1][w] = ((unsigned long long)(unsigned)vmax_w << 32) | (unsigned)(~rmin_w);
            unsigned *dst = &s_ring[((j - 1) / kBqRing) & 1][(j - 1) % kBqRing][w * PPL];
#pragma unroll
            for (int i = 0; i < PPL; i++) dst[i] = hb[i];
        }
        __syncthreads();
        unsigned long long key = s_key[j & 1][0];
#pragma unroll
        for (int q = 1; q < W; q++) { const unsigned long long o = s_key[j & 1][q]; key = o > key ? o : key; }
        const int vmax = (int)(key >> 32);
        if ((j % kBqRing) == 0 || vmax == 0) emit_batch(j);
        if (vmax == 0) { done = true; break; }
        cur = brev_n<L>((int)~(unsigned)key);
        if (tid == 0) sidx[j] = cur;
    }
    if (!done) {                                         // the last sample has had no round of its own
        const float4 c = sp[cur];
        unsigned hb[PPL];
#pragma unroll
        for (int i = 0; i < PPL; i++)
            hb[i] = __ballot_sync(0xffffffffu, sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z)) < r2);
        if (lane == 0) {
            unsigned *dst = &s_ring[((m - 1) / kBqRing) & 1][(m - 1) % kBqRing][w * PPL];
#pragma unroll
            for (int i = 0; i < PPL; i++) dst[i] = hb[i];
        }
        __syncthreads();
        emit_batch(m);
    }
    __syncthreads();
    if (done) {                                          // samples j .. m-1 are point 0 = sample 0: their balls are row 0
        for (int q = j + tid; q < m; q += W * 32) sidx[q] = 0;
        // dup_rows == 0: the caller only walks the distinct centroids (live_cnt), rounded up to a tile of four
        const int lim = dup_rows ? m : min(m, (j + 3) & ~3);
        for (int q = j + w; q < lim; q += W)
            for (int l = lane; l < K; l += 32) ball_idx[(size_t)q * K + l] = s_row0[l];
        __syncthreads();
    }
    for (int q = tid; q < m; q += W * 32) idxs[q] = sidx[q];
    for (int e = tid; e < m * 3; e += W * 32) {
        const int q = e / 3, cc = e - q * 3;
        const float4 v = sp[sidx[q]];
        new_xyz[e] = cc == 0 ? v.x : (cc == 1 ? v.y : v.z);
    }
    // number of DISTINCT samples: centroids live .. m-1 repeat centroid 0 (same ball, same features)
    if (live_cnt && tid == 0) live_cnt[cloud] = done ? j : m;
}

// order[0 .. b) = the clouds sorted by weight, heaviest first (counting sort over weights clamped to [0, 1023]; the
// order inside a weight class is whatever the atomics give - it only shapes the schedule, never a result).  One CTA.
__global__ void __launch_bounds__(1024)
order_clouds_kernel(const int *__restrict__ weight, int b, int *__restrict__ order) {
    __shared__ int s_hist[1024];
    __shared__ int s_warp[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    s_hist[tid] = 0;
    __syncthreads();
    for (int c = tid; c < b; c += 1024) {
        const int wv = min(max(__ldg(weight + c), 0), 1023);
        atomicAdd(&s_hist[1023 - wv], 1);                // bucket 0 = heaviest
    }
    __syncthreads();
    const int h = s_hist[tid];
    int incl = h;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += u;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = 0;
    for (int q = 0; q < warp; q++) before += s_warp[q];
    __syncthreads();
    s_hist[tid] = before + incl - h;                     // exclusive start of the bucket
    __syncthreads();
    for (int c = tid; c < b; c += 1024) {
        const int wv = min(max(__ldg(weight + c), 0), 1023);
        order[atomicAdd(&s_hist[1023 - wv], 1)] = c;
    }
}

}  // namespace mgar
