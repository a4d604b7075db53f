// Set-abstraction shared MLP on the 5th-generation tensor cores (tcgen05 + TMEM), fused with the
// grouping gather and the max-pool: the one dense contraction of the LiDAR hot path.
//
// Reference: StackSAModuleMSG.forward (pointnet2_stack/pointnet2_modules.py:78-112) after
// QueryAndGroup (pointnet2_utils.py:123-159).  Shape handled here: 2 layers, C0 = 3 + C_in <= 8,
// C1 = 64, C2 = 128, nsample = 32, max-pool (BASELINE configs[1]: 4 -> 64 -> 128); anything else
// runs on the exact-fp32 kernel in sa_mlp.cu.
//
// Layer 1 (K = C0, a few FMAs per output) stays on the CUDA cores in exact fp32.  Layer 2 is
//     D^T[c, pos] = sum_k W2[c, k] * h1[pos, k]        (M = 128 channels, N = 128 positions, K = 64)
// as tcgen05.mma kind::tf32 with fp32 accumulation in TMEM.  To keep fp32-level accuracy (the
// parity bar is 1e-5) every operand is split v = hi + lo with hi = tf32(v) and the product is
// evaluated as lo*hi + hi*lo + hi*hi (3xTF32: the dropped lo*lo term is ~2^-22 relative).
// Because the output channel is the M (TMEM lane) dimension, a thread of the epilogue owns one
// channel and the 32 samples of a centroid are 32 consecutive TMEM columns: BatchNorm, ReLU and
// the max over samples are a register loop with no shuffles, and the only HBM write is out[m, c].
//
// One persistent CTA per SM, warp-specialised:
//   warps 0-3   epilogue: tcgen05.ld 32x32b.x32 -> scale/shift, max, ReLU -> coalesced store
//   warp  4     TMEM allocation + layer-2 MMA issue (24 MMAs per 128-position tile); warp 13: layer-1 MMA issue
//   warps 5-12  producers (two groups of four, one smem stage each): gather (xyz - centre | feat),
//               layer 1, hi/lo split, 128B-swizzled K-major store of the B tile
// mbarriers: full/empty per smem stage, tfull/tempty per TMEM accumulator stage (4 x 128 columns).
#include <mutex>
#include <vector>
#include <stdlib.h>

#include <type_traits>

#include "sa_mlp_tc.cuh"
#include "stack_prefix.cuh"
#include "tc_ptx.cuh"

#ifndef MGAR_TC_TERMS
#define MGAR_TC_TERMS 3   // 3xTF32; timing experiments build with fewer
#endif

namespace mgar {

constexpr int kTcThreads = 15 * 32;
constexpr int kTcStages = 3;                  // shared-memory stages of the h1 operand
constexpr int kAccStages = 2;                 // TMEM stages of the layer-2 accumulator (128 columns each)
// TMEM columns: layer-2 accumulators 0,128; layer-1 accumulators 256,320 (64 each); W2 operand hi 384, lo 448
constexpr uint32_t kTmD1 = 256, kTmW2Hi = 384, kTmW2Lo = 448;
constexpr uint32_t kXBytes = 128 * 8 * 4;       // one 128 x 8 fp32 layer-1 input tile (no swizzle): 4 KB
constexpr uint32_t kW1Bytes = 64 * 8 * 4;       // W1^T operand, 64 x 8: 2 KB
constexpr uint32_t kOpBytes = 128 * 64 * 4;     // one 128 x 64 fp32 operand tile: 32 KB
constexpr uint32_t kTcSmem = kTcStages * 2 * kOpBytes /* h1 hi,lo */ + 2 * 2 * kXBytes /* x hi,lo x2 */ + 2 * kW1Bytes + 1024;


// byte offset of element (row r, 16-byte chunk c16 of the 64-wide K extent) inside an operand tile
__device__ __forceinline__ uint32_t tile_off(int r, int c16) {
    const int kb = c16 >> 3, cw = c16 & 7;
    return (uint32_t)(kb * 16384 + (r >> 3) * 1024 + (r & 7) * 128 + ((cw ^ (r & 7)) << 4));
}


template <int C0>
__global__ void __launch_bounds__(kTcThreads, 1)
sa_mlp_tc_kernel(const __grid_constant__ TcLayer1 L1, const __grid_constant__ TcArgs A) {
    static_assert(C0 >= 3 && C0 <= 7, "layer-1 K is padded to 8 with one column for the folded bias");
    extern __shared__ uint8_t smem_raw[];
    // full[3] empty[3] | tfull[2] tempty[2] | xfull[2] xempty[2] | dfull[2] dempty[2]
    __shared__ __align__(8) uint64_t s_bar[18];
    __shared__ uint32_t s_tmem;

    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B atoms need 1024 B alignment
    const uint32_t b_stage0 = base;                                 // h1 stage s: hi at +s*2*kOpBytes, lo at +kOpBytes more
    const uint32_t x_stage0 = base + kTcStages * 2 * kOpBytes;      // x stage: hi at +xs*2*kXBytes, lo at +kXBytes more
    const uint32_t w1_hi = x_stage0 + 2 * 2 * kXBytes, w1_lo = w1_hi + kW1Bytes;
    const uint32_t bar0 = smem_u32(s_bar);
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (3 + s); };
    auto TFULL = [&](int t) { return bar0 + 8u * (6 + t); };
    auto TEMPTY = [&](int t) { return bar0 + 8u * (8 + t); };
    auto XFULL = [&](int x) { return bar0 + 8u * (10 + x); };
    auto XEMPTY = [&](int x) { return bar0 + 8u * (12 + x); };
    auto DFULL = [&](int x) { return bar0 + 8u * (14 + x); };
    auto DEMPTY = [&](int x) { return bar0 + 8u * (16 + x); };

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef MGAR_TC_TRACE
#define TRACE(slot, i) do { if (blockIdx.x == 0 && lane == 0 && (i) >= 100 && (i) < 132) A.trace[((i) - 100) * 16 + (slot)] = clock64(); } while (0)
#else
#define TRACE(slot, i) do { } while (0)
#endif
    const int M = A.M;
    const int n_tiles = (M + 3) >> 2;
    // contiguous tile range per CTA: consecutive tiles belong to the same person, so the running
    // group max of the epilogue is flushed only a handful of times
    const int t_beg = (int)((long long)blockIdx.x * n_tiles / gridDim.x);
    const int t_end = (int)((long long)(blockIdx.x + 1) * n_tiles / gridDim.x);
    const int my_tiles = t_end - t_beg;

    // ---- one-time setup: barriers, TMEM, weight operands ---------------------------------------------
    if (tid == 0) {
        for (int s = 0; s < kTcStages; s++) { mbar_init(FULL(s), 256); mbar_init(EMPTY(s), 1); }
        for (int t = 0; t < kAccStages; t++) { mbar_init(TFULL(t), 1); mbar_init(TEMPTY(t), 128); }
        for (int x = 0; x < 2; x++) {
            mbar_init(XFULL(x), 128); mbar_init(XEMPTY(x), 1);
            mbar_init(DFULL(x), 1); mbar_init(DEMPTY(x), 256);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // W1^T (64 channels x K=8: C0 folded weights, the folded BatchNorm shift in column C0, zeros) as the
    // B operand of the layer-1 MMA, hi/lo split, no-swizzle K-major
    if (tid >= 160 && tid < 224) {
        const int ch = tid - 160;
        float k0[4], k1[4];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float v = k < C0 ? L1.w[k][ch] : (k == C0 ? L1.sh[ch] : 0.f);
            if (k < 4) k0[k] = v; else k1[k - 4] = v;
        }
        sts_split4(w1_hi + k8_off(ch, 0), w1_lo + k8_off(ch, 0), k0);
        sts_split4(w1_hi + k8_off(ch, 1), w1_lo + k8_off(ch, 1), k1);
        fence_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    // W2 lives in TENSOR MEMORY as the A operand of the layer-2 MMA (row = output channel = lane,
    // column = k): the MMA then fetches only the h1 tile from shared memory, and shared memory is left to
    // three h1 stages.
    if (warp < 4) {
        const int c = tid;
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll
        for (int kh = 0; kh < 2; kh++) {
            uint32_t hi[32], lo[32];
#pragma unroll
            for (int k = 0; k < 32; k++) {
                float h, l;
                split_tf32(__ldg(A.w2t + (size_t)(kh * 32 + k) * kTcC2 + c), h, l);
                hi[k] = __float_as_uint(h); lo[k] = __float_as_uint(l);
            }
            tc_st32(lane_base + kTmW2Hi + kh * 32, hi);
            tc_st32(lane_base + kTmW2Lo + kh * 32, lo);
        }
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (warp < 4) {
        // ===== epilogue: thread = output channel, 32 consecutive TMEM columns = one centroid ==========
        // This role is a serial chain per tile (wait, TMEM load, reduce, store), so it is kept free of long
        // dependent sequences: four independent partial maxima per centroid, no integer division (the
        // group / person boundary is tracked incrementally), running output pointer.
        const int c = tid;
        const float scale = __ldg(A.sc2 + c), shift = __ldg(A.sh2 + c);
        const int gs = A.group_size;
        int m_next = t_beg * 4;                                   // first centroid of the next tile
        int grp = A.gmax ? m_next / gs : 0, g_end = (grp + 1) * gs;
        float gm = 0.f;
        float *outp = A.out ? A.out + (size_t)m_next * A.ld_out + c : nullptr;
        auto flush = [&]() {
            if (!(A.group_zero && __ldg(A.group_zero + grp)))
                atomicMax(reinterpret_cast<int *>(A.gmax + (size_t)grp * kTcC2 + c), __float_as_int(gm));
        };
        auto reduce32 = [&](const uint32_t (&v)[32]) {
            float a[4];
#pragma unroll
            for (int q = 0; q < 4; q++) a[q] = __fmaf_rn(__uint_as_float(v[q]), scale, shift);
#pragma unroll
            for (int s = 4; s < 32; s++) a[s & 3] = fmaxf(a[s & 3], __fmaf_rn(__uint_as_float(v[s]), scale, shift));
            return fmaxf(fmaxf(fmaxf(a[0], a[1]), fmaxf(a[2], a[3])), 0.f);   // ReLU commutes with the max
        };
        for (int i = 0; i < my_tiles; i++) {
            const int t = i % kAccStages, u = i / kAccStages;
            if (warp == 0) TRACE(14, i);
            mbar_wait(TFULL(t), u & 1);
            if (warp == 0) TRACE(0, i);
            tc_fence_after();
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(t * 128);
            float res[4];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                uint32_t v0[32], v1[32];
                tc_ld32(taddr + (2 * h) * 32, v0);
                tc_ld32(taddr + (2 * h + 1) * 32, v1);
                tc_wait_ld();
                res[2 * h] = reduce32(v0);
                res[2 * h + 1] = reduce32(v1);
            }
            tc_fence_before();
            mbar_arrive_relaxed(TEMPTY(t));
            if (warp == 0) TRACE(1, i);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int m = m_next + q;
                if (m < M) {
                    if (outp) outp[(size_t)q * A.ld_out] = res[q];
                    if (A.gmax) {
                        if (m >= g_end) { flush(); grp++; g_end += gs; gm = 0.f; }
                        gm = fmaxf(gm, res[q]);
                    }
                }
            }
            m_next += 4;
            if (outp) outp += (size_t)4 * A.ld_out;
        }
        if (A.gmax && my_tiles > 0) flush();
    } else if (warp == 4 || warp == 14) {
        // ===== layer-2 MMA issuers: warp 4 takes the even tiles, warp 14 the odd ones (different TMEM
        // accumulators and smem stages), so while one of them is between tiles - commits, barrier round
        // trips - the other keeps the tensor pipe fed.
        // ===== MMA issuer.  Per tile: layer 1 = 3 x (128x64x8), layer 2 = 24 x (128x128x8) =================
        // A 128x128x8 tf32 MMA is only 64 tensor-pipe cycles, so the ISSUE path must stay well under that
        // per instruction: the whole warp runs the (warp-uniform) control flow so descriptors are built on
        // the uniform datapath, only the tcgen05 instructions are predicated on one elected lane, and a
        // descriptor step is a single add on its low word.
        const uint32_t idesc2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        for (int i = (warp == 4 ? 0 : 1); i < my_tiles; i += 2) {
            const int s = i % kTcStages, j = i / kTcStages, t = i % kAccStages, u = i / kAccStages;
            TRACE(5, i);
            mbar_wait(FULL(s), j & 1);
            TRACE(6, i);
            mbar_wait(TEMPTY(t), (u & 1) ^ 1);
            TRACE(7, i);
            tc_fence_after();
            const uint32_t b_hi = b_stage0 + (uint32_t)s * 2 * kOpBytes;
            const uint64_t bd_hi = make_desc(b_hi), bd_lo = make_desc(b_hi + kOpBytes);
            const uint32_t d = tmem + (uint32_t)(t * 128);
            if (elect_one()) {
                // small terms first: W2lo*h1hi, W2hi*h1lo, then W2hi*h1hi.  K step kk: 8 TMEM columns of W2,
                // 32 bytes inside the 128B-swizzled row (second K block 16 KB further) of h1.
#pragma unroll
                for (int term = 0; term < MGAR_TC_TERMS; term++) {
                    const uint32_t a0 = tmem + (term == 0 ? kTmW2Lo : kTmW2Hi);
                    const uint64_t b0 = term == 1 ? bd_lo : bd_hi;
#pragma unroll
                    for (int kk = 0; kk < 8; kk++) {
                        const uint32_t koff = (uint32_t)((kk >> 2) * 16384 + (kk & 3) * 32);
                        tc_mma_tf32_ts(d, a0 + (uint32_t)(kk * 8), b0 + (uint64_t)(koff >> 4), idesc2, (term | kk) ? 1u : 0u);
                    }
                }
                tc_commit(EMPTY(s));     // smem stage reusable once these MMAs have read it
                tc_commit(TFULL(t));     // accumulator ready for the epilogue
            }
            __syncwarp();
            TRACE(8, i);
        }
    } else if (warp == 13) {
        // ===== layer-1 MMA issuer: its own warp, so these three small MMAs (and their barrier round trips)
        // are off the serial path of the layer-2 issuer; the tensor pipe interleaves the two streams.
        const uint32_t idesc1 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint64_t w1d_hi = make_desc_k8(w1_hi), w1d_lo = make_desc_k8(w1_lo);
        for (int i = 0; i < my_tiles; i++) {
            // D1[pos, ch] = x[pos, 0:8] . W1^T[ch, 0:8] (the folded bias rides in column C0)
            const int xs = i & 1, jx = i >> 1;
            mbar_wait(XFULL(xs), jx & 1);
            mbar_wait(DEMPTY(xs), (jx & 1) ^ 1);
            tc_fence_after();
            const uint32_t x_hi = x_stage0 + (uint32_t)xs * 2 * kXBytes;
            const uint64_t xd_hi = make_desc_k8(x_hi), xd_lo = make_desc_k8(x_hi + kXBytes);
            const uint32_t d1 = tmem + kTmD1 + (uint32_t)(xs * 64);
            if (elect_one()) {
                tc_mma_tf32(d1, xd_lo, w1d_hi, idesc1, 0);
                tc_mma_tf32(d1, xd_hi, w1d_lo, idesc1, 1);
                tc_mma_tf32(d1, xd_hi, w1d_hi, idesc1, 1);
                tc_commit(XEMPTY(xs));
                tc_commit(DFULL(xs));
            }
            __syncwarp();
        }
    } else {
        // ===== gather + converters (8 warps) =================================================================
        // warp = (position block pw = TMEM lane quarter, channel half).  Per tile:
        //   (a) the 4 warps of half 0 gather (xyz - centre | feat), split hi/lo and store the 128 x 8 layer-1
        //       input tile (bias column = 1) one tile AHEAD; the gather itself (idx -> point, two dependent
        //       loads) is software-pipelined two more tiles ahead in registers;
        //   (b) all 8 warps read their position's 32 layer-1 outputs back from TMEM (tcgen05.ld), apply
        //       ReLU, split hi/lo and store the 128B-swizzled K-major h1 operand of the layer-2 MMA.
        const int pw = warp & 3, half = (warp - 5) >> 2;
        const int r = pw * 32 + lane;

        int u_cloud = A.uniform ? (t_beg * 4 + pw) / A.m_u : 0, u_end = (u_cloud + 1) * A.m_u;   // one division, then incremental
        auto load_idx = [&](int i, int &iv, int &em) {          // local tile i -> this lane's sample index
            const int m = (t_beg + i) * 4 + pw;
            iv = -1; em = 0;
            if (i < my_tiles && m < M) {
                iv = __ldg(A.idx + (size_t)m * kTcNS + lane);
                if (A.empty_mask) em = A.empty_mask[m];
            }
        };
        // raw gathered row (before the centre is subtracted) + centre; live = centroid exists and ball not empty
        auto load_pt = [&](int i, int iv, int em, float (&v)[C0], float (&q)[3], bool &live) {
            const int m = (t_beg + i) * 4 + pw;
            const int first = __shfl_sync(0xffffffffu, iv, 0);
            live = (i < my_tiles) && (m < M) && first >= 0 && !em;
            if (live) {
                long long start;
                if (A.uniform) {
                    while (m >= u_end) { u_cloud++; u_end += A.m_u; }     // tiles arrive in increasing order
                    start = (long long)u_cloud * A.n_u;
                } else {
                    start = __ldg(A.pre_xyz + find_batch_tc(A.pre_new, A.B, m));
                }
                const long long p = start + (iv < 0 ? 0 : iv);
                if (C0 == 4 && A.packed4) {
                    const float4 t4 = __ldg(reinterpret_cast<const float4 *>(A.xyz + p * 4));
                    v[0] = t4.x; v[1] = t4.y; v[2] = t4.z; v[C0 - 1] = t4.w;
                } else {
                    v[0] = __ldg(A.xyz + p * A.xyz_stride + 0);
                    v[1] = __ldg(A.xyz + p * A.xyz_stride + 1);
                    v[2] = __ldg(A.xyz + p * A.xyz_stride + 2);
#pragma unroll
                    for (int cc = 3; cc < C0; cc++) v[cc] = __ldg(A.features + p * A.feat_stride + (cc - 3));
                }
                q[0] = __ldg(A.new_xyz + (size_t)m * 3 + 0);
                q[1] = __ldg(A.new_xyz + (size_t)m * 3 + 1);
                q[2] = __ldg(A.new_xyz + (size_t)m * 3 + 2);
            }
        };
        // layer-1 input row of local tile i -> x stage i & 1
        auto store_x = [&](int i, const float (&v)[C0], const float (&q)[3], bool live) {
            const int xs = i & 1, jx = i >> 1;
            float k0[4] = {0.f, 0.f, 0.f, 0.f}, k1[4] = {0.f, 0.f, 0.f, 0.f};
            if (live) {
                k0[0] = __fsub_rn(v[0], q[0]); k0[1] = __fsub_rn(v[1], q[1]); k0[2] = __fsub_rn(v[2], q[2]);
#pragma unroll
                for (int cc = 3; cc < C0; cc++) { if (cc < 4) k0[cc] = v[cc]; else k1[cc - 4] = v[cc]; }
            }
            if (C0 < 4) k0[C0] = 1.f; else k1[C0 - 4] = 1.f;   // multiplies the folded BatchNorm shift
            const uint32_t x_hi = x_stage0 + (uint32_t)xs * 2 * kXBytes, x_lo = x_hi + kXBytes;
            mbar_wait(XEMPTY(xs), (jx & 1) ^ 1);
            sts_split4(x_hi + k8_off(r, 0), x_lo + k8_off(r, 0), k0);
            sts_split4(x_hi + k8_off(r, 1), x_lo + k8_off(r, 1), k1);
            fence_async_smem();
            mbar_arrive(XFULL(xs));
        };

        int iv_a = -1, em_a = 0, iv_b = -1, em_b = 0;
        float vb[C0], qb[3];
        bool live_b = false;
        if (half == 0) {
            load_idx(0, iv_b, em_b);
            load_pt(0, iv_b, em_b, vb, qb, live_b);
            load_idx(1, iv_b, em_b);
            load_idx(2, iv_a, em_a);
            if (my_tiles > 0) store_x(0, vb, qb, live_b);
            load_pt(1, iv_b, em_b, vb, qb, live_b);
        }
        for (int i = 0; i < my_tiles; i++) {
            if (half == 0 && i + 1 < my_tiles) {
                store_x(i + 1, vb, qb, live_b);                 // gathered during the previous iteration
                iv_b = iv_a; em_b = em_a;
                load_pt(i + 2, iv_b, em_b, vb, qb, live_b);     // points two tiles ahead
                load_idx(i + 3, iv_a, em_a);                    // indices three tiles ahead
            }
            const int xs = i & 1, jx = i >> 1, s = i % kTcStages, j = i / kTcStages;
            if (warp == 8) TRACE(9, i);
            mbar_wait(DFULL(xs), jx & 1);
            if (warp == 8) TRACE(10, i);
            tc_fence_after();
            uint32_t h[32];
#ifndef MGAR_TC_NOLDTM
            tc_ld32(tmem + ((uint32_t)(pw * 32) << 16) + kTmD1 + (uint32_t)(xs * 64 + half * 32), h);
            tc_wait_ld();
#else
            for (int q = 0; q < 32; q++) h[q] = 0x3f800000u + q + i;
#endif
            tc_fence_before();
            mbar_arrive_relaxed(DEMPTY(xs));

            const uint32_t b_hi = b_stage0 + (uint32_t)s * 2 * kOpBytes, b_lo = b_hi + kOpBytes;
            if (warp == 8) TRACE(11, i);
            mbar_wait(EMPTY(s), (j & 1) ^ 1);
            if (warp == 8) TRACE(12, i);
#pragma unroll
            for (int c8 = 0; c8 < 8; c8++) {
                float y[4];
#pragma unroll
                for (int q = 0; q < 4; q++) y[q] = fmaxf(__uint_as_float(h[c8 * 4 + q]), 0.f);
                const uint32_t off = tile_off(r, half * 8 + c8);
                sts_split4(b_hi + off, b_lo + off, y);
            }
            fence_async_smem();          // generic-proxy stores -> visible to the tensor core (async proxy)
            mbar_arrive(FULL(s));
            if (warp == 8) TRACE(13, i);
        }
    }

    // ---- teardown ------------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

}  // namespace mgar

using namespace mgar;

// Workspace of the live-person / live-tile walk: one counter + the list (the caller owns it: SURVEY 8b).
extern "C" size_t mgar_sa_mlp_tc_workspace_bytes(int M, int group_size) {
    if (M <= 0 || group_size <= 0) return 0;
    const size_t groups = ((size_t)M + group_size - 1) / group_size;
    // the longest list is the live-unit one: up to four octets per centroid, a person's padded to a multiple of 16
    const size_t upt = (size_t)tc_units_per_tile();
    const size_t units = groups * (((size_t)tc_units_per_row() * group_size + upt - 1) / upt * upt);
    return (4 + units) * sizeof(int);
}

extern "C" int mgar_sa_mlp_tc_unit_slots() { return kTcNS / tc_units_per_row(); }

extern "C" int mgar_sa_mlp_tc_supported(int nsample, int C_in, int n_layers, const int *channels_host, int pool_avg) {
    return nsample == kTcNS && C_in >= 0 && C_in <= 4 && n_layers == 2 && channels_host && channels_host[0] == 3 + C_in &&
           channels_host[1] == kTcC1 && channels_host[2] == kTcC2 && !pool_avg;
}

// nsample 16 / 64 when ONLY the per-group max is wanted (out == NULL): the 3xFP16 kernel walks blocks of 32 positions
// (TcArgs::ns).  MGAR_TC_NS32_ONLY=1 switches it off (A/B runs against the fp32 kernel).
extern "C" int mgar_sa_mlp_tc_group_max_supported(int nsample, int C_in, int n_layers, const int *channels_host, int pool_avg) {
    if (nsample == kTcNS) return mgar_sa_mlp_tc_supported(nsample, C_in, n_layers, channels_host, pool_avg);
    if (nsample != 16 && nsample != 64) return 0;
    if (getenv("MGAR_TC_NS32_ONLY")) return 0;
    const char *mode = getenv("MGAR_TC_MODE");
    if (mode && mode[0] == 't') return 0;
    return mgar_sa_mlp_tc_supported(kTcNS, C_in, n_layers, channels_host, pool_avg);
}

extern "C" int mgar_sa_mlp_fwd_tc(int B, int M, int nsample, int C_in, const float *xyz, int xyz_stride,
                                  const int *xyz_batch_cnt, const float *new_xyz, const int *new_xyz_batch_cnt,
                                  int uniform_n, int uniform_m, const float *features, int feat_stride, const int *idx,
                                  const unsigned char *empty_mask, int n_layers, const int *channels_host,
                                  const float *params, const float *params_host, float *out, int ld_out,
                                  float *group_max, int group_size, const int *group_zero, const int *group_live_cnt,
                                  const int *group_point_cnt, int pool_avg, void *workspace, size_t workspace_bytes,
                                  void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0, "sa_mlp_tc: bad size");
    const bool ns32 = nsample == kTcNS;
    MGAR_REQUIRE(ns32 ? mgar_sa_mlp_tc_supported(nsample, C_in, n_layers, channels_host, pool_avg)
                      : (!out && group_max && mgar_sa_mlp_tc_group_max_supported(nsample, C_in, n_layers, channels_host, pool_avg)),
                 "sa_mlp_tc: only nsample=32 (16 / 64 with out == NULL and a group_max), C_in<=4, channels [3+C_in,64,128], "
                 "max-pool run on the tensor-core path");
    if (M == 0) return MGAR_OK;
    MGAR_REQUIRE(xyz && new_xyz && idx && params && params_host && (out || group_max), "sa_mlp_tc: null pointer");
    MGAR_REQUIRE(features || C_in == 0, "sa_mlp_tc: features missing");
    MGAR_REQUIRE(xyz_stride >= 3 && (C_in == 0 || feat_stride >= C_in), "sa_mlp_tc: bad row stride");
    MGAR_REQUIRE(!group_max || group_size >= 1, "sa_mlp_tc: group_size must be >= 1");
    const bool uniform = uniform_n > 0;
    MGAR_REQUIRE(uniform || (xyz_batch_cnt && new_xyz_batch_cnt), "sa_mlp_tc: batch counts missing");
    // nsample != 32: the kernel's unit is a block of 32 positions - Mv blocks, gsv per group, muv per cloud
    MGAR_REQUIRE(ns32 || (uniform && ((long long)group_size * nsample) % 32 == 0 && ((long long)uniform_m * nsample) % 32 == 0 &&
                          M % group_size == 0),
                 "sa_mlp_tc: nsample %d needs the uniform layout with whole 32-position blocks per group", nsample);
    const int Mv = ns32 ? M : (int)((long long)M * nsample / 32);
    const int gsv = ns32 ? group_size : group_size * nsample / 32, muv = ns32 ? uniform_m : uniform_m * nsample / 32;
    const int C0 = 3 + C_in;
    TcLayer1 L1;
    for (int o = 0; o < kTcC1; o++) {
        const float sc = params_host[C0 * kTcC1 + o];
        for (int c = 0; c < 8; c++) L1.w[c][o] = c < C0 ? params_host[c * kTcC1 + o] * sc : 0.f;   // BN scale folded
        L1.sh[o] = params_host[C0 * kTcC1 + kTcC1 + o];
    }
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (!uniform) {
        if (int rc = ws.bind(workspace, workspace_bytes, B, "sa_mlp_fwd_tc")) return rc;
        if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    }
    TcArgs A;
    A.B = B; A.M = Mv; A.C_in = C_in; A.ns = nsample;
    A.xyz = xyz; A.xyz_stride = xyz_stride; A.new_xyz = new_xyz;
    A.features = features; A.feat_stride = feat_stride;
    A.pre_xyz = ws.b; A.pre_new = ws.a; A.uniform = uniform ? 1 : 0; A.n_u = uniform_n; A.m_u = muv;
    A.idx = idx; A.empty_mask = empty_mask;
    A.w2t = params + C0 * kTcC1 + 2 * kTcC1;
    A.sc2 = A.w2t + kTcC1 * kTcC2; A.sh2 = A.sc2 + kTcC2;
    A.out = out; A.ld_out = ld_out;
    A.gmax = group_max; A.group_size = gsv > 0 ? gsv : 1; A.group_zero = group_zero;
    A.tr_zext = nullptr; A.tr_arg = nullptr; A.tr_sum = nullptr; A.tr_sq = nullptr;
    A.l1_dev = nullptr;
    A.live_groups = nullptr; A.n_live = nullptr; A.live_tiles = nullptr; A.live_units = nullptr; A.m_shift = -1;
    A.packed4 = (C0 == 4 && xyz_stride == 4 && feat_stride == 4 && features == xyz + 3 &&
                 (reinterpret_cast<uintptr_t>(xyz) & 15) == 0) ? 1 : 0;
#ifdef MGAR_TC_TRACE
    static long long *d_trace = nullptr;
    if (!d_trace) MGAR_CUDA(cudaMalloc(&d_trace, 32 * 16 * sizeof(long long)));
    MGAR_CUDA(cudaMemsetAsync(d_trace, 0, 32 * 16 * sizeof(long long), st));
    A.trace = d_trace;
#endif
    if (group_max) {
        const size_t groups = ((size_t)Mv + A.group_size - 1) / A.group_size;
        MGAR_CUDA(cudaMemsetAsync(group_max, 0, groups * kTcC2 * sizeof(float), st));
    }
    const int n_tiles = (Mv + 3) / 4;
    const int grid = n_tiles < kNumSMs ? n_tiles : kNumSMs;     // persistent: one CTA per SM
#ifndef MGAR_TC_TRACE
    {   // default: the fp16-split kernel (same accuracy, half the tensor cycles); MGAR_TC_MODE=tf32 keeps 3xTF32
        const char *mode = getenv("MGAR_TC_MODE");
        if (!(mode && mode[0] == 't')) {
            // Persons flagged empty only produce a zero row: when the per-centroid features are not stored, walk the
            // live persons only (a list compacted on the device - no host round trip)
            // Asked for with ld_out == -1 (see the header); MGAR_TC_SKIP_EMPTY=0/1 overrides for A/B runs.
            const char *skip = getenv("MGAR_TC_SKIP_EMPTY");
            const bool want = skip ? skip[0] == '1' : ld_out == -1;
            if (want && C0 == 4 && !out && group_max && group_zero && uniform && uniform_m == group_size && gsv % 4 == 0 &&
                M % group_size == 0) {
                const int G = M / group_size;
                int *live = static_cast<int *>(workspace);
                if (!live || workspace_bytes < mgar_sa_mlp_tc_workspace_bytes(M, group_size)) {
                    set_error("sa_mlp_tc: the live-person walk needs a workspace of mgar_sa_mlp_tc_workspace_bytes(M, group_size) "
                              "= %zu bytes, got %zu", mgar_sa_mlp_tc_workspace_bytes(M, group_size), workspace ? workspace_bytes : (size_t)0);
                    return MGAR_ENOMEM;
                }
                if (group_live_cnt && group_point_cnt && ns32 && M < (1 << 28) && !getenv("MGAR_TC_NO_LIVE_TILES") &&
                    !getenv("MGAR_TC_NO_LIVE_UNITS")) {
                    // distinct samples of the distinct centroids only (see TcArgs::live_units)
                    if (int rc = tc_build_live_units(group_zero, group_live_cnt, group_point_cnt, idx, G, gsv, uniform_n, live + 4,
                                                     live, st)) return rc;
                    A.live_units = live + 4; A.n_live = live;
                    for (int sft = 0; sft < 31; sft++) if ((1 << sft) == gsv) A.m_shift = sft;
                } else if (group_live_cnt && !getenv("MGAR_TC_NO_LIVE_TILES")) {
                    // distinct centroids only (see TcArgs::live_tiles): the sampling kernel reported how many each person has
                    if (int rc = tc_build_live_tiles(group_zero, group_live_cnt, G, gsv, nsample, live + 4, live, st)) return rc;
                    A.live_tiles = live + 4; A.n_live = live;
                    for (int sft = 0; sft < 31; sft++) if ((1 << sft) == gsv) A.m_shift = sft;
                } else {
                    if (int rc = tc_compact_live_groups(group_zero, G, live + 4, live, st)) return rc;
                    A.live_groups = live + 4; A.n_live = live;
                }
            }
            return sa_mlp_tc16_launch(L1, A, C0, grid, st);
        }
    }
#endif
#ifdef MGAR_TC_TRACE
#define MGAR_TC_TRACE_DUMP                                                                      \
    {                                                                                           \
        static long long h[32 * 16];                                                            \
        cudaStreamSynchronize(st);                                                              \
        cudaMemcpy(h, d_trace, sizeof(h), cudaMemcpyDeviceToHost);                              \
        for (int i = 0; i < 32; i++) {                                                          \
            fprintf(stderr, "tctrace %d", 100 + i);                                             \
            for (int q = 0; q < 15; q++) fprintf(stderr, " %lld", h[i * 16 + q] ? h[i * 16 + q] - h[5] : -1); \
            fprintf(stderr, "\n");                                                              \
        }                                                                                       \
    }
#else
#define MGAR_TC_TRACE_DUMP
#endif
    MGAR_REQUIRE(ns32, "sa_mlp_tc: MGAR_TC_MODE=tf32 runs nsample 32 only");
#define MGAR_TC_CASE(c0)                                                                                        \
    if (C0 == c0) {                                                                                             \
        MGAR_CUDA(cudaFuncSetAttribute(sa_mlp_tc_kernel<c0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmem)); \
        sa_mlp_tc_kernel<c0><<<grid, kTcThreads, kTcSmem, st>>>(L1, A);                                         \
        count_launch();                                                                                         \
        MGAR_TC_TRACE_DUMP                                                                                      \
        return check_launch("sa_mlp_tc");                                                                       \
    }
    MGAR_TC_CASE(3) MGAR_TC_CASE(4) MGAR_TC_CASE(5) MGAR_TC_CASE(6) MGAR_TC_CASE(7)
#undef MGAR_TC_CASE
    set_error("sa_mlp_tc: unsupported C0=%d", C0);
    return MGAR_EINVAL;
}
