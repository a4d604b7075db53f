// FPS + ball query for WRAPPED 512-slot clouds: only the distinct 32-slot blocks take part in a round.
//
// Per-person pooling fills the slots k >= c of a person that pooled c < 512 points with slot k mod c
// (roipoint_pool3d_kernel.cu:86-97), so a pooled cloud is periodic: slot k holds point k mod p.  The reference's sampling
// kernel (sampling_gpu.cu:25-140) and ball query (ball_query_gpu.cu:16-66) nevertheless visit all 512 slots in every one
// of the p rounds.  fps_bq_cta_kernel (fps_bq_cta.cuh) did the same.  Here a round touches only the first
// nb = ceil(p / 32) blocks of 32 slots - every other slot is a copy of one of them, with the SAME running minimum by
// construction - and the two places where the copies matter are reconstructed exactly:
//
//   * tie order.  Among equal maxima the reference's winner minimises rank(k) = brev9(k) over ALL 512 slots (header of
//     fps.cu).  The set of maximal slots is a union of copy classes {q, q + p, q + 2p, ...}, so the winner is the smallest
//     rk(s) = min over the class of slot s of brev9(k), over the maximal processed slots s; rk(s) is a per-slot constant
//     computed once per cloud.  The winning INDEX brev9(rk) may be a copy slot >= 32 nb - it is what the reference returns.
//   * ball rows.  The first K hits in index order over all 512 slots: blocks < nb have real ballots; the hit mask of a
//     block s >= nb is read out of those (bit i = hit of slot (32 s + i) mod p: two funnel shifts) when the rows are
//     assembled - every kBqRing rounds, off the round's critical path as in fps_bq_cta_kernel.
//
// The period is a HINT that is verified: after staging, the CTA checks slot k == slot k - p for every k >= p (one compare
// per slot, once per cloud) and falls back to p = 512 (all 16 blocks, the plain algorithm) when it does not hold, so the
// result never depends on the hint.  Two warps per cloud: block b belongs to warp b & 1, register slot b >> 1; the round
// loop is instantiated for PE = ceil(nb / 2) = 1 .. 8 register slots per lane.
#pragma once
#include "fps_bq_cta.cuh"

namespace mgar {

constexpr int kDupN = 512, kDupBlocks = 16, kDupW = 2;

struct DupCtx {
    const float4 *sp; int *sidx; int *rows; int *s_row0;
    unsigned (*ring)[16]; unsigned (*key)[2][2];
    int *idxs; float *new_xyz; int *ball_idx; int *live_cnt;
    float r2; int K, m, cloud, dup_rows, p, nb;
};

// 32 hit bits starting at bit `pos` of the processed masks M[0 .. nb) (ring row `rw`; block b sits at (b & 1) * 8 + (b >> 1))
__device__ __forceinline__ unsigned dup_get32(const unsigned *rw, int nb, int pos) {
    const int wi = pos >> 5, sh = pos & 31;
    const unsigned a = wi < nb ? rw[(wi & 1) * 8 + (wi >> 1)] : 0u;
    const unsigned b = wi + 1 < nb ? rw[((wi + 1) & 1) * 8 + ((wi + 1) >> 1)] : 0u;
    return __funnelshift_r(a, b, sh);
}

template <int PE>
__device__ __forceinline__ void bq_dup_body(const DupCtx &C) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int K = C.K, m = C.m, p = C.p, nb = C.nb;
    const int lane0 = (int)bq_pin(lane == 0 ? 1u : 0u), thread0 = (int)bq_pin(tid == 0 ? 1u : 0u);
    const uint32_t sp_a = bq_pin(bq_smem_u32(C.sp)), sidx_a = bq_pin(bq_smem_u32(C.sidx));
    const uint32_t key_w = bq_pin(bq_smem_u32(&C.key[0][0][0]) + (uint32_t)w * 8u);
    const uint32_t key_a = bq_pin(bq_smem_u32(&C.key[0][0][0]));
    const uint32_t ring_a = bq_pin(bq_smem_u32(&C.ring[0][0]) + (uint32_t)(w * 8 * 4));
    const float r2 = C.r2;

    float px[PE], py[PE], pz[PE], t[PE];
    int rk[PE];
    bool act[PE];
#pragma unroll
    for (int i = 0; i < PE; i++) {
        const int blk = 2 * i + w, k = blk * 32 + lane;
        act[i] = blk < nb;                                   // warp-uniform
        const float4 q = C.sp[act[i] ? k : 0];
        // a slot without a block sits 1e18 away (its squared distance, ~3e36, is finite and beyond any radius: never a
        // hit) with a running minimum of -1 (fminf keeps it; distances are >= 0, so it never wins): the round needs no
        // per-slot activity test
        px[i] = act[i] ? q.x : 1e18f; py[i] = act[i] ? q.y : 1e18f; pz[i] = act[i] ? q.z : 1e18f;
        t[i] = act[i] ? 1e10f : -1.f;
        int best = INT_MAX;
        if (act[i])
            for (int kk = k % p; kk < kDupN; kk += p) best = min(best, brev_n<9>(kk));     // once per cloud
        rk[i] = best;
    }
    // block s of the emitter: where its 32 bits start inside the period (constant per cloud)
    const int es = lane & 15;
    const int e_start = es >= nb ? (32 * es) % p : 0;

    int c_emitted = 0;
    auto emit_batch = [&](int c_end) {
        const int c0 = c_emitted;
        const int half = lane >> 4, s = es;
        int *row = C.rows + (w * 2 + half) * K;
        for (int rr = 2 * w; rr < kBqRing && c0 + rr < c_end; rr += 2 * kDupW) {
            const int c = c0 + rr + half;
            const bool valid = c < c_end;
            const unsigned *rw = C.ring[c & (2 * kBqRing - 1)];
            unsigned mask = 0u;
            if (valid) {
                if (s < nb) {
                    mask = rw[(s & 1) * 8 + (s >> 1)];
                } else {                                     // a block of copies: slot 32 s + i holds point (32 s + i) mod p
                    const int L = 32 * nb, n_low = min(32, L - e_start);
                    mask = dup_get32(rw, nb, e_start);
                    if (n_low < 32) {
                        mask &= (1u << n_low) - 1u;
                        mask |= dup_get32(rw, nb, e_start + n_low - p) << n_low;
                    }
                }
            }
            const int pc = __popc(mask);
            int incl = pc;
#pragma unroll
            for (int o = 1; o < 16; o *= 2) {
                const int up = __shfl_up_sync(0xffffffffu, incl, o, 16);
                if (s >= o) incl += up;
            }
            int pos = incl - pc;
            while (mask && pos < K) {                        // this block's hits, in index (= bit) order
                row[pos++] = s * 32 + __ffs(mask) - 1;
                mask &= mask - 1;
            }
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int ch = c0 + rr + h;
                if (ch >= c_end) break;
                const int cnt = __shfl_sync(0xffffffffu, incl, h * 16 + 15);
                const int *rh = C.rows + (w * 2 + h) * K;
                const int first = rh[0];
                for (int l = lane; l < K; l += 32) {
                    const int v = cnt == 0 ? (l == 0 ? -1 : 0) : (l < cnt ? rh[l] : first);
                    C.ball_idx[(size_t)ch * K + l] = v;
                    if (ch == 0) C.s_row0[l] = v;
                }
            }
            __syncwarp();
        }
        c_emitted = c_end;
    };

    auto store_masks = [&](int c, const unsigned (&hb)[PE]) {
        unsigned h8[8];
#pragma unroll
        for (int i = 0; i < 8; i++) h8[i] = i < PE ? hb[i] : 0u;
        const uint32_t a = ring_a + (uint32_t)(c & (2 * kBqRing - 1)) * 64u;
        bq_sts_u4_if(lane0, a, h8[0], h8[1], h8[2], h8[3]);
        if (PE > 4) bq_sts_u4_if(lane0, a + 16, h8[4], h8[5], h8[6], h8[7]);
    };

    int cur = 0, j = 1;
    bool done = false;
    if (tid == 0) C.sidx[0] = 0;
    for (; j < m; j++) {
        const float4 c = bq_lds_f4(sp_a + (uint32_t)cur * 16u);
        unsigned hb[PE];
#pragma unroll
        for (int i = 0; i < PE; i++) {
            const float d = sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z));
            hb[i] = __ballot_sync(0xffffffffu, d < r2);
            t[i] = fminf(d, t[i]);
        }
        float mx = t[0];
#pragma unroll
        for (int i = 1; i < PE; i++) mx = fmaxf(mx, t[i]);
        const int vb = __float_as_int(mx);                   // warp w = 1 of an odd nb = 1 cloud holds only -1: never the maximum
        const int vmax_w = __reduce_max_sync(0xffffffffu, vb);
        const float vmax_f = __int_as_float(vmax_w);
        int cand = INT_MAX;
#pragma unroll
        for (int i = 0; i < PE; i++) cand = t[i] == vmax_f ? min(cand, rk[i]) : cand;
        const int rmin_w = __reduce_min_sync(0xffffffffu, cand);
        store_masks(j - 1, hb);
        const uint32_t ka = key_a + (uint32_t)(j & 1) * 16u;
        bq_sts_u2_if(lane0, key_w + (uint32_t)(j & 1) * 16u, ~(unsigned)rmin_w, (unsigned)vmax_w);
        __syncthreads();
        unsigned kk[4];
        bq_lds_u4(ka, kk);
        // (value << 32 | ~rank) per warp as one unsigned 64-bit key, compared as signed on the value word: -1 (0xbf800000)
        // of a warp without active blocks is a negative int and loses against any distance >= 0
        const bool take1 = ((int)kk[3] > (int)kk[1]) || (kk[3] == kk[1] && kk[2] > kk[0]);
        const unsigned vmax = take1 ? kk[3] : kk[1], nrank = take1 ? kk[2] : kk[0];
        if (((j & (kBqRing - 1)) == 0) | (vmax == 0)) emit_batch(j);
        if (vmax == 0) { done = true; break; }
        cur = brev_n<9>((int)~nrank);
        bq_sts_u1_if(thread0, sidx_a + (uint32_t)j * 4u, (unsigned)cur);
    }
    if (!done) {                                             // the last sample has had no round of its own
        const float4 c = C.sp[cur];
        unsigned hb[PE];
#pragma unroll
        for (int i = 0; i < PE; i++) {
            hb[i] = __ballot_sync(0xffffffffu, sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z)) < r2);
        }
        store_masks(m - 1, hb);
        __syncthreads();
        emit_batch(m);
    }
    __syncthreads();
    if (done) {
        for (int q = j + tid; q < m; q += kDupW * 32) C.sidx[q] = 0;
        const int lim = C.dup_rows ? m : min(m, (j + 7) & ~7);
        for (int q = j + w; q < lim; q += kDupW)
            for (int l = lane; l < K; l += 32) C.ball_idx[(size_t)q * K + l] = C.s_row0[l];
        __syncthreads();
    }
    for (int q = tid; q < m; q += kDupW * 32) C.idxs[q] = C.sidx[q];
    for (int q = tid; q < m; q += kDupW * 32) {
        const float4 v = C.sp[C.sidx[q]];
        C.new_xyz[q * 3 + 0] = v.x; C.new_xyz[q * 3 + 1] = v.y; C.new_xyz[q * 3 + 2] = v.z;
    }
    if (C.live_cnt && tid == 0) C.live_cnt[C.cloud] = done ? j : m;
}

// One CTA of two warps per cloud of 512 slots.  period[cloud] (optional) = the caller's claim "slot k holds point k mod p".
// (93 registers; capping them for 12 / 14 / 16 resident CTAs per SM measured 0.284 / 0.290 / 0.306 ms against 0.268 for the stage)
__global__ void __launch_bounds__(kDupW * 32)
fps_bq_dup_kernel(const float *__restrict__ xyz, int xs, int *__restrict__ idxs, float *__restrict__ new_xyz,
                  int *__restrict__ ball_idx, float radius, int K, int b, int m, const int *__restrict__ order,
                  const int *__restrict__ period, int *__restrict__ live_cnt, int dup_rows,
                  const int *__restrict__ n_split = nullptr) {
    extern __shared__ float4 bq_sm4[];
    __shared__ __align__(16) unsigned s_key[2][2][2];
    __shared__ __align__(16) unsigned s_ring[2 * kBqRing][16];
    __shared__ int s_row0[64];
    __shared__ __align__(8) unsigned long long s_mbar;
    float4 *sp = bq_sm4;
    int *sidx = reinterpret_cast<int *>(sp + kDupN);
    const int tid = threadIdx.x;
    // n_split: the first *n_split entries of `order` (the long clouds) belong to fps_bq_cta_kernel<4, 4> running beside this launch
    const int slot = (int)blockIdx.x + (n_split ? __ldg(n_split) : 0);
    if (slot >= b) return;
    const int cloud = order ? __ldg(order + slot) : slot;
    if (cloud < 0 || cloud >= b) return;
    xyz += (size_t)cloud * kDupN * xs;
    if (xs == 4 && (reinterpret_cast<uintptr_t>(xyz) & 15) == 0) {
        const uint32_t bar = bq_smem_u32(&s_mbar);
        if (tid == 0) bq_mbar_init(bar, 1);
        __syncthreads();
        if (tid == 0) bq_bulk_load(bq_smem_u32(sp), xyz, (uint32_t)(kDupN * sizeof(float4)), bar);
        bq_mbar_wait(bar, 0);
    } else {
        for (int k = tid; k < kDupN; k += kDupW * 32)
            sp[k] = make_float4(xyz[(size_t)k * xs], xyz[(size_t)k * xs + 1], xyz[(size_t)k * xs + 2], 0.f);
        __syncthreads();
    }
    // the period is a hint: verified here, p = 512 (every block) when the cloud is not the wrap it was said to be
    int p = period ? __ldg(period + cloud) : kDupN;
    p = p < 1 ? 1 : (p > kDupN ? kDupN : p);
    int bad = 0;
    for (int k = p + tid; k < kDupN; k += kDupW * 32) {
        const float4 a = sp[k], c = sp[k - p];
        bad |= (__float_as_int(a.x) != __float_as_int(c.x)) | (__float_as_int(a.y) != __float_as_int(c.y)) |
               (__float_as_int(a.z) != __float_as_int(c.z));
    }
    if (__syncthreads_or(bad)) p = kDupN;
    DupCtx C;
    C.sp = sp; C.sidx = sidx; C.rows = sidx + m; C.s_row0 = s_row0; C.ring = s_ring; C.key = s_key;
    C.idxs = idxs + (size_t)cloud * m; C.new_xyz = new_xyz + (size_t)cloud * m * 3; C.ball_idx = ball_idx + (size_t)cloud * m * K;
    C.live_cnt = live_cnt; C.r2 = __fmul_rn(radius, radius); C.K = K; C.m = m; C.cloud = cloud; C.dup_rows = dup_rows;
    C.p = p; C.nb = (p + 31) >> 5;
    switch ((C.nb + 1) >> 1) {
        case 1: bq_dup_body<1>(C); break;
        case 2: bq_dup_body<2>(C); break;
        case 3: bq_dup_body<3>(C); break;
        case 4: bq_dup_body<4>(C); break;
        case 5: bq_dup_body<5>(C); break;
        case 6: bq_dup_body<6>(C); break;
        case 7: bq_dup_body<7>(C); break;
        default: bq_dup_body<8>(C); break;
    }
}

}  // namespace mgar
