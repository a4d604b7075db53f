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
//     into a shared-memory ring; every kBqRing rounds the W warps turn the ring into rows - two centroids per warp pass
//     (half-warp per centroid: 16 lanes = the 16 blocks of a cloud), so the serial append loop and the row store are
//     off the round's critical path and amortised.
//   * clouds are taken in the caller's `order` (longest first: the pooled point count is known before the sampling),
//     so the short clouds fill in behind the long ones instead of the long ones defining a tail.
//
// The kernel is bound by the ALU pipe (FSETP / VOTE / FMNMX / SEL issue at half rate; ncu: 71 % ALU-pipe active,
// "math pipe throttle" the top stall), so the round is written to a minimum of instructions: shared-memory addresses
// live in registers (explicit ld/st.shared), stores by one lane are predicated instead of branched, the (value, rank)
// exchange is one 64-bit word per warp.  W = 2, PPL = 8 is the default for 512-point clouds: half the cross-warp
// overhead of W = 4 per point, while a round's chain stays ~4x shorter than the one-warp kernel's.
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

constexpr int kBqRing = 8;      // rounds per emission batch (2 centroids per warp pass)

template <int BITS>
__device__ __forceinline__ int brev_n(int v) { return BITS == 0 ? 0 : (int)(__brev((unsigned)v) >> (32 - BITS)); }

template <int V> struct CtLog2 { static constexpr int v = 1 + CtLog2<V / 2>::v; };
template <> struct CtLog2<1> { static constexpr int v = 0; };

__device__ __forceinline__ uint32_t bq_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float4 bq_lds_f4(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void bq_lds_u4(uint32_t a, unsigned (&v)[4]) {
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(a));
}
__device__ __forceinline__ void bq_lds_u2(uint32_t a, unsigned (&v)[2]) {
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v[0]), "=r"(v[1]) : "r"(a));
}
// stores executed by the lanes whose `on` is non-zero (predicated: no branch, no reconvergence point)
__device__ __forceinline__ void bq_sts_u2_if(int on, uint32_t a, unsigned x, unsigned y) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %0, 0;\n\t@p st.shared.v2.u32 [%1], {%2, %3};\n\t}" ::"r"(on), "r"(a), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void bq_sts_u4_if(int on, uint32_t a, unsigned x, unsigned y, unsigned z, unsigned w) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %0, 0;\n\t@p st.shared.v4.u32 [%1], {%2, %3, %4, %5};\n\t}" ::"r"(on), "r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ void bq_sts_u1_if(int on, uint32_t a, unsigned x) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %0, 0;\n\t@p st.shared.u32 [%1], %2;\n\t}" ::"r"(on), "r"(a), "r"(x) : "memory");
}

// a value the compiler must keep in a register (it otherwise re-derives shared-memory addresses and lane predicates
// from S2R / LEA chains in every round, ~20 instructions of a ~140-instruction round)
__device__ __forceinline__ uint32_t bq_pin(uint32_t v) {
    uint32_t r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}

// ---- bulk asynchronous copy (the TMA engine's 1-D form): one instruction moves a cloud's tile global -> shared -------
__device__ __forceinline__ void bq_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bq_bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bq_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}

template <int PPL>
__device__ __forceinline__ void bq_store_masks(int on, uint32_t a, const unsigned (&hb)[PPL]) {
    if constexpr (PPL >= 4) {
#pragma unroll
        for (int i = 0; i < PPL; i += 4) bq_sts_u4_if(on, a + 4 * i, hb[i], hb[i + 1], hb[i + 2], hb[i + 3]);
    } else if constexpr (PPL == 2) {
        bq_sts_u2_if(on, a, hb[0], hb[1]);
    } else {
        bq_sts_u1_if(on, a, hb[0]);
    }
}

// kRedux: the W (value, ~rank) keys are combined by two warp reductions over lanes that each hold one key instead of a
// compare tree over all W keys in every lane: ~9 fewer instructions per warp and round at ~25 cycles more latency - for
// the loaded case (many clouds per SM: a round costs its instruction count times the resident warps), not for a single
// frame (few clouds: a round costs its latency).
template <int W, int PPL, bool kRedux = false>
__global__ void __launch_bounds__(W * 32)
fps_bq_cta_kernel(const float *__restrict__ xyz, int xs, int *__restrict__ idxs, float *__restrict__ new_xyz,
                  int *__restrict__ ball_idx, float radius, int K, int b, int m, const int *__restrict__ order,
                  int *__restrict__ live_cnt, int dup_rows, const int *__restrict__ n_split = nullptr) {
    constexpr int NBLK = W * PPL, NB = CtLog2<NBLK>::v, n = 32 * NBLK, L = 5 + NB;
    static_assert(NBLK <= 16 && (NBLK & (NBLK - 1)) == 0, "a cloud is at most 16 blocks of 32 points");
    extern __shared__ float4 bq_sm4[];
    __shared__ __align__(16) unsigned s_key[2][W > 1 ? W : 2][2];    // per parity and warp: (max distance bits, ~rank)
    __shared__ __align__(16) unsigned s_ring[2 * kBqRing][16];       // hit masks of the last 16 centroids
    __shared__ int s_row0[64];
    __shared__ __align__(8) unsigned long long s_mbar;
    __shared__ __align__(8) unsigned s_pin[W * 32][2];
    float4 *sp = bq_sm4;                                   // [n] staged cloud
    int *sidx = reinterpret_cast<int *>(sp + n);           // [m] sampled indices
    int *rows = sidx + m;                                  // [W][2][K] rows being assembled

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // n_split: this launch takes only the first *n_split entries of `order` (the long clouds; the others belong to the
    // periodic-cloud kernel running beside it, fps_bq_dup.cuh)
    if (n_split && (int)blockIdx.x >= __ldg(n_split)) return;
    const int cloud = order ? __ldg(order + blockIdx.x) : (int)blockIdx.x;
    if (cloud < 0 || cloud >= b) return;
    xyz += (size_t)cloud * n * xs; idxs += (size_t)cloud * m;
    new_xyz += (size_t)cloud * m * 3;
    ball_idx += (size_t)cloud * m * K;

    if (xs == 4 && (reinterpret_cast<uintptr_t>(xyz) & 15) == 0) {
        // the cloud is one contiguous tile of n x 16 bytes (the xyz | feature rows of the pooled array): staged by ONE
        // bulk asynchronous copy (cp.async.bulk, the TMA engine) that signals an mbarrier - no per-thread LDG + STS
        const uint32_t bar = bq_smem_u32(&s_mbar);
        if (tid == 0) bq_mbar_init(bar, 1);
        __syncthreads();
        if (tid == 0) bq_bulk_load(bq_smem_u32(sp), xyz, (uint32_t)(n * sizeof(float4)), bar);
        bq_mbar_wait(bar, 0);
    } else {
        for (int k = tid; k < n; k += W * 32)
            sp[k] = make_float4(xyz[(size_t)k * xs], xyz[(size_t)k * xs + 1], xyz[(size_t)k * xs + 2], 0.f);
        __syncthreads();
    }

    // Per-thread loop constants go THROUGH shared memory once: ptxas sees through `mov` pins and rematerialises such values
    // inside the round (r^2 from the constant bank + an FMUL, the rank from S2R / BREV / SHF / LOP3: 8 of ~100 instructions
    // per round); a value loaded from shared memory it has to keep.
    float r2 = __fmul_rn(radius, radius);
    int rank0 = (brev_n<5>(lane) << NB) | (w * PPL);
    {
        s_pin[tid][0] = (unsigned)rank0; s_pin[tid][1] = __float_as_uint(r2);
        __syncwarp();
        unsigned pv[2];
        bq_lds_u2(bq_smem_u32(&s_pin[tid][0]), pv);
        rank0 = (int)pv[0]; r2 = __uint_as_float(pv[1]);
    }
    const int lane0 = (int)bq_pin(lane == 0 ? 1u : 0u), thread0 = (int)bq_pin(tid == 0 ? 1u : 0u);
    const uint32_t sp_a = bq_pin(bq_smem_u32(sp)), sidx_a = bq_pin(bq_smem_u32(sidx));
    const uint32_t key_a = bq_pin(bq_smem_u32(&s_key[0][0][0]));
    const uint32_t key_w = bq_pin(key_a + (uint32_t)w * 8u);
    const uint32_t ring_a = bq_pin(bq_smem_u32(&s_ring[0][0]) + (uint32_t)(w * PPL * 4));
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
        const int c0 = c_emitted;
        const int half = lane >> 4, s = lane & 15;
        int *row = rows + (w * 2 + half) * K;
        for (int rr = 2 * w; rr < kBqRing && c0 + rr < c_end; rr += 2 * W) {
            const int c = c0 + rr + half;
            const bool valid = c < c_end;
            unsigned mask = (valid && s < NBLK) ? s_ring[c & (2 * kBqRing - 1)][brev_n<NB>(s)] : 0u;
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
        const float4 c = bq_lds_f4(sp_a + (uint32_t)cur * 16u);
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
        bq_store_masks<PPL>(lane0, ring_a + (uint32_t)((j - 1) & (2 * kBqRing - 1)) * 64u, hb);
        unsigned vmax, nrank;                              // the cloud's maximum and ~(rank of the winner)
        if constexpr (W == 1) {
            vmax = (unsigned)vmax_w; nrank = ~(unsigned)rmin_w;
        } else {
            const uint32_t ka = key_a + (uint32_t)(j & 1) * (uint32_t)(W * 8);
            bq_sts_u2_if(lane0, key_w + (uint32_t)(j & 1) * (uint32_t)(W * 8), ~(unsigned)rmin_w, (unsigned)vmax_w);
            __syncthreads();
            if constexpr (kRedux) {
                unsigned kk[2];                            // lane l holds the key of warp l mod W: kk[0] = ~rank, kk[1] = value
                bq_lds_u2(ka + (uint32_t)(lane & (W - 1)) * 8u, kk);
                vmax = __reduce_max_sync(0xffffffffu, kk[1]);
                nrank = __reduce_max_sync(0xffffffffu, kk[1] == vmax ? kk[0] : 0u);
            } else {
            // (value << 32 | ~rank) as one unsigned 64-bit key per warp: the largest key wins
            unsigned long long key = 0;
#pragma unroll
            for (int q = 0; q < W; q += 2) {
                unsigned kk[4];
                bq_lds_u4(ka + (uint32_t)q * 8u, kk);
                const unsigned long long k0 = ((unsigned long long)kk[1] << 32) | kk[0];
                const unsigned long long k1 = ((unsigned long long)kk[3] << 32) | kk[2];
                const unsigned long long kq = k1 > k0 ? k1 : k0;
                key = kq > key ? kq : key;
            }
            vmax = (unsigned)(key >> 32); nrank = (unsigned)key;
            }
        }
        if (((j & (kBqRing - 1)) == 0) | (vmax == 0)) {
            if constexpr (W == 1) __syncwarp();
            emit_batch(j);
        }
        if (vmax == 0) { done = true; break; }
        cur = brev_n<L>((int)~nrank);
        bq_sts_u1_if(thread0, sidx_a + (uint32_t)j * 4u, (unsigned)cur);
    }
    if (!done) {                                         // the last sample has had no round of its own
        const float4 c = sp[cur];
        unsigned hb[PPL];
#pragma unroll
        for (int i = 0; i < PPL; i++)
            hb[i] = __ballot_sync(0xffffffffu, sqdist(__fsub_rn(px[i], c.x), __fsub_rn(py[i], c.y), __fsub_rn(pz[i], c.z)) < r2);
        bq_store_masks<PPL>(lane0, ring_a + (uint32_t)((m - 1) & (2 * kBqRing - 1)) * 64u, hb);
        __syncthreads();
        emit_batch(m);
    }
    __syncthreads();
    if (done) {                                          // samples j .. m-1 are point 0 = sample 0: their balls are row 0
        for (int q = j + tid; q < m; q += W * 32) sidx[q] = 0;
        // dup_rows == 0: the caller only walks the distinct centroids (live_cnt), rounded up to eight
        // (a tile of the MLP kernel = 128 positions = 4 centroids of 32 samples, 8 of 16)
        const int lim = dup_rows ? m : min(m, (j + 7) & ~7);
        for (int q = j + w; q < lim; q += W)
            for (int l = lane; l < K; l += 32) ball_idx[(size_t)q * K + l] = s_row0[l];
        __syncthreads();
    }
    for (int q = tid; q < m; q += W * 32) idxs[q] = sidx[q];
    for (int q = tid; q < m; q += W * 32) {
        const float4 v = sp[sidx[q]];
        new_xyz[q * 3 + 0] = v.x; new_xyz[q * 3 + 1] = v.y; new_xyz[q * 3 + 2] = v.z;
    }
    // number of DISTINCT samples: centroids live .. m-1 repeat centroid 0 (same ball, same features)
    if (live_cnt && tid == 0) live_cnt[cloud] = done ? j : m;
}

// order[0 .. b) = the clouds sorted by weight, heaviest first (counting sort over weights clamped to [0, 1023]; the
// order inside a weight class is whatever the atomics give - it only shapes the schedule, never a result).  One CTA.
// n_heavy (optional) = number of clouds with weight >= heavy_from: they are the first n_heavy entries of order.
__global__ void __launch_bounds__(1024)
order_clouds_kernel(const int *__restrict__ weight, int b, int *__restrict__ order, int *__restrict__ n_heavy = nullptr,
                    int heavy_from = 0) {
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
    // bucket t holds weight 1023 - t: the heavy clouds are the buckets 0 .. 1023 - heavy_from
    if (n_heavy && tid == min(max(1023 - heavy_from, 0), 1023)) *n_heavy = heavy_from > 1023 ? 0 : before + incl;
    __syncthreads();
    for (int c = tid; c < b; c += 1024) {
        const int wv = min(max(__ldg(weight + c), 0), 1023);
        order[atomicAdd(&s_hist[1023 - wv], 1)] = c;
    }
}

}  // namespace mgar
