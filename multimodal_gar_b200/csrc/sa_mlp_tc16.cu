// Set-abstraction shared MLP on tcgen05 with FP16 operand splitting ("3xFP16"): same contract and the same
// fp32-level accuracy as sa_mlp_tc.cu (3xTF32), at twice the tensor rate.
//
// Reference: StackSAModuleMSG.forward (pointnet2_stack/pointnet2_modules.py:78-112) after
// QueryAndGroup (pointnet2_utils.py:123-159); shape [3 + C_in, 64, 128], nsample 32, max-pool.
//
// The 3xTF32 kernel runs with the tensor pipe 97 % busy (profiles/r01_full_summary_3xtf32.md), so the only
// way down is fewer tensor cycles per tile.  kind::f16 moves 128x128x16 per instruction in the time
// kind::tf32 moves 128x128x8, and an fp16 significand (11 bits) is as wide as a tf32 one, so
//     v = hi + lo,  hi = fp16(v), lo = fp16(v - hi)        W.h ~ Wlo.hhi + Whi.hlo + Whi.hhi
// carries the same 22 bits per operand as the tf32 split: 12 MMAs of 128x128x16 per tile instead of 24 of
// 128x128x8.  What fp16 lacks is exponent range, so both operands are moved into it with EXACT
// power-of-two scales that are undone in the epilogue:
//   * W2 row c (an output channel) is multiplied by 2^ec so that its largest entry sits in [2^13, 2^14), and by
//     sign(scale2[c]) so that BatchNorm's affine is non-decreasing in the accumulator: the epilogue then takes
//     the max over the 32 samples FIRST (one FMNMX per value) and applies scale/shift once per centroid.
//     fl(a*max(D)+b) == max(fl(a*D+b)) for a >= 0 because rounding is monotone, so nothing changes numerically.
//   * every centroid (32 positions = one gather warp) gets 2^-e from an a-priori bound of its layer-1 outputs,
//     |z1[k]| <= sum_c wmax[c]*|x[c]| + shmax, reduced over the warp with one REDUX; the layer-1 input rows
//     (and the bias column) are multiplied by it, so h1 leaves layer 1 already inside [0, 2^15).  Entries
//     more than 2^18 below the bound lose bits only below 2^-40 of it.
// Layer 1 itself stays 3xTF32 (K = 8, three small MMAs).
//
// Warp roles, barriers and the software pipeline are those of sa_mlp_tc.cu:
//   warps 0-3 epilogue | 4-7 / 8-11 gather + convert of the even / odd tiles (a tile's whole serial chain belongs to one
//   group of four warps) | 12, 13 layer-2 issue (even / odd tiles) | 14 layer-1 issue
#include <cuda_fp16.h>
#include <stdlib.h>

#include "sa_mlp_tc.cuh"
#include "stack_prefix.cuh"
#include "tc_ptx.cuh"

namespace mgar {

namespace {

// converter groups (four warps each; group g owns the tiles i = g mod kGroups) and layer-2 issuing warps
#ifndef MGAR_TC16_GROUPS
#define MGAR_TC16_GROUPS 2
#endif
#ifndef MGAR_TC16_ISSUERS
#define MGAR_TC16_ISSUERS 2
#endif
// Measured on the bench step (MLP stage, ms per 64 frames): groups x issuers 2x2 0.408 | 2x1 0.455 | 3x1 0.465 | 3x2 0.409 (112
// registers, a few spilled words): a third converter group buys nothing - the stage is not paced by the converters' own
// instruction count - while a second issuing warp is worth 10 %.
constexpr int kGroups = MGAR_TC16_GROUPS, kIssuers = MGAR_TC16_ISSUERS;
// an issuer must always meet the same accumulator stage: with kIssuers not dividing kAcc two issuers would wait on one
// mbarrier for different phases, and a parity wait cannot tell phases two apart (this hangs)
static_assert(kIssuers == 1 || kIssuers == 2, "layer-2 issuers: 1 or 2 (kAcc = 2 accumulator stages)");
constexpr int kWarpL2 = 4 + 4 * kGroups, kWarpL1 = kWarpL2 + kIssuers;      // warps 0-3 epilogue, 4 .. kWarpL2-1 converters
constexpr int kThreads = (kWarpL1 + 1) * 32;
constexpr int kStages = 4;                      // shared-memory stages of the h1 operand (hi + lo = 32 KB each)
constexpr int kAcc = 2;                         // TMEM stages of the layer-2 accumulator (128 columns each)
// TMEM columns: layer-2 accumulators 0,128; W2 fp16 hi 256, lo 288 (32 each); layer-1 accumulators 320 + 64 g (64 each)
constexpr uint32_t kTmW2Hi = 256, kTmW2Lo = 288, kTmD1 = 320;
// MGAR_TC16_XTMEM=1: the layer-1 input tile travels registers -> TENSOR MEMORY (tcgen05.st, the A operand of the
// layer-1 MMA is read from TMEM) instead of registers -> shared memory -> async-proxy fence: 16 columns per stage behind D1
// Measured: no gain (MLP stage 0.410 against 0.407 ms) - the default stays shared memory.
#ifndef MGAR_TC16_XTMEM
#define MGAR_TC16_XTMEM 0
#endif
constexpr bool kXTmem = MGAR_TC16_XTMEM != 0;
// MGAR_TC16_DEFER_X=1: the x tile's async-proxy fence + XFULL arrive wait until the group has read the previous D1 (the
// stores have long completed by then).  Measured: no gain either (0.4045 against 0.4025 ms).
#ifndef MGAR_TC16_DEFER_X
#define MGAR_TC16_DEFER_X 0
#endif
constexpr bool kDeferX = MGAR_TC16_DEFER_X != 0;
constexpr uint32_t kTmX = kTmD1 + 64 * kGroups;
static_assert(kTmX + (kXTmem ? 16 * kGroups : 0) <= 512, "TMEM holds 512 columns");
constexpr uint32_t kXBytes = 128 * 8 * 4;       // one 128 x 8 fp32 layer-1 input tile (no swizzle): 4 KB
constexpr uint32_t kW1Bytes = 64 * 8 * 4;       // W1^T operand, 64 x 8: 2 KB
constexpr uint32_t kHBytes = 128 * 64 * 2;      // one 128 x 64 fp16 operand tile (one 128-byte swizzled row per position): 16 KB
constexpr uint32_t kSmem = kStages * 2 * kHBytes + kGroups * 2 * kXBytes + 2 * kW1Bytes + 1024;   // (the x stages are unused with kXTmem)
constexpr int kRing = 16;
// distinct-sample walk: a list entry covers kUnitSlots consecutive sample slots of a centroid's ball row, a tile holds
// kUnitsPerTile = 128 / kUnitSlots of them (TcArgs::live_units).  8 slots: tiles x 0.824 of the distinct-centroid walk on the
// bench data, 4 slots: x 0.793 - and the same stage time (0.396 ms either way: twice the list entries and row probes, up
// to eight centroids per gather warp), so 8 stays.
#ifndef MGAR_TC16_UNIT_SLOTS
#define MGAR_TC16_UNIT_SLOTS 8
#endif
constexpr int kUnitSlots = MGAR_TC16_UNIT_SLOTS, kUnitsPerTile = 128 / kUnitSlots, kUnitsPerRow = kTcNS / kUnitSlots;
constexpr int kUnitShift = kUnitSlots == 4 ? 3 : 2;          // log2(kUnitsPerRow): entry = centroid << kUnitShift | unit
static_assert(kUnitSlots == 4 || kUnitSlots == 8, "units of 4 or 8 sample slots");                       // per-centroid scale ring (gather runs at most ~9 tiles ahead of the epilogue)

__device__ __forceinline__ void tc_mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {   // fp16 operands, A read from TMEM
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
                 "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
                 : "memory");
}

// ReLU(a), ReLU(b) -> packed fp16 pairs hi, lo (a in the low half) with hi + lo = ReLU(.) to 22 bits.
// hi is rounded TOWARD ZERO so that the remainder of a positive value is >= 0; cvt's .relu then clamps exactly
// the negative inputs in both conversions (a < 0: hi = 0, remainder = a < 0, lo = 0) and no FMNMX is needed.
__device__ __forceinline__ void split_relu_h2(float a, float b, uint32_t &hi, uint32_t &lo) {
    asm("cvt.rz.relu.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(b), "f"(a));
    const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(__fsub_rn(b, f.y)), "f"(__fsub_rn(a, f.x)));
}
// (a, b) -> packed fp16 pair of the rounded values (a in the low half) and of the remainders
__device__ __forceinline__ void split_h2(float a, float b, uint32_t &hi, uint32_t &lo) {
    const __half2 h = __floats2half2_rn(a, b);
    const float2 f = __half22float2(h);
    const __half2 l = __floats2half2_rn(__fsub_rn(a, f.x), __fsub_rn(b, f.y));
    hi = *reinterpret_cast<const uint32_t *>(&h);
    lo = *reinterpret_cast<const uint32_t *>(&l);
}

// byte offset of 16-byte chunk c16 (8 fp16 = channels 8*c16 .. 8*c16+7) of position row r in an h1 tile
__device__ __forceinline__ uint32_t h_off(int r, int c16) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4));
}

__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(a, fmaxf(b, c)); }   // one FMNMX3
// max of 32 values in 16 three-input operations (two independent chains)
__device__ __forceinline__ float max32(const uint32_t (&v)[32]) {
    float a = max3(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]));
    float b = max3(__uint_as_float(v[3]), __uint_as_float(v[4]), __uint_as_float(v[5]));
#pragma unroll
    for (int s = 6; s < 30; s += 4) {
        a = max3(a, __uint_as_float(v[s]), __uint_as_float(v[s + 1]));
        b = max3(b, __uint_as_float(v[s + 2]), __uint_as_float(v[s + 3]));
    }
    return max3(a, b, fmaxf(__uint_as_float(v[30]), __uint_as_float(v[31])));
}

// training-statistics epilogue: max, first index attaining it, sum and sum of squares of 32 accumulator values.
// The 3xFP16 accumulator carries ~2^-22 of the products' magnitude, so two DIFFERENT samples closer than that can
// swap places against an fp32 evaluation; `near` reports a runner-up within 2^-14 of the centroid's rms (a wide
// margin over the split's error) and the caller re-evaluates those few entries in exact fp32.
__device__ __forceinline__ void tr_reduce32(const uint32_t (&v)[32], float &mx, int &arg, float &s, float &q, bool &near) {
    mx = max32(v);
    arg = 31;
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
    float r0 = -INFINITY, r1 = -INFINITY;               // runner-up: largest value different from the maximum
#pragma unroll
    for (int k = 30; k >= 0; k--) arg = __uint_as_float(v[k]) == mx ? k : arg;
#pragma unroll
    for (int k = 0; k < 32; k += 2) {
        const float a = __uint_as_float(v[k]), b = __uint_as_float(v[k + 1]);
        s0 += a; s1 += b;
        q0 = __fmaf_rn(a, a, q0); q1 = __fmaf_rn(b, b, q1);
        r0 = fmaxf(r0, a == mx ? -INFINITY : a);
        r1 = fmaxf(r1, b == mx ? -INFINITY : b);
    }
    s = s0 + s1; q = q0 + q1;
    near = mx - fmaxf(r0, r1) <= 6.1035156e-5f * sqrtf(q * 0.03125f);
}

// MGAR_TC16_EXP=n: timing-only experiment builds (results WRONG by construction; never the shipped library):
//   1 no h1 stores (converters keep the TMEM read and the split)   2 eight instead of twelve layer-2 MMAs
//   3 no x-tile stores   4 one instead of two layer-1 MMAs
#ifndef MGAR_TC16_EXP
#define MGAR_TC16_EXP 0
#endif
#ifdef MGAR_TC16_TRACE
__device__ long long g_tc16_trace[32 * 16];        // timeline of CTA 0, local tiles 100..131 (debug build only)
#define TR(slot, i) do { if (blockIdx.x == 0 && lane == 0 && (i) >= 100 && (i) < 132) g_tc16_trace[((i) - 100) * 16 + (slot)] = clock64(); } while (0)
#else
#define TR(slot, i) do { } while (0)
#endif

// 2-MMA packing of a 5-column 3xTF32 product (see kPack2 below).  v = columns 0..3, v4 = column 4; hi = tf32
// truncation, lo = exact remainder, as sts_split4.  a0/a1 = the two 16-byte K-halves of the row in tile A, b0/b1 in B.
__device__ __forceinline__ void split5(const float (&v)[4], float v4, float (&hi)[5], float (&lo)[5]) {
#pragma unroll
    for (int q = 0; q < 5; q++) {
        const float x = q < 4 ? v[q] : v4;
        hi[q] = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
        lo[q] = __fsub_rn(x, hi[q]);
    }
}
__device__ __forceinline__ void sts4(uint32_t a, float x, float y, float z, float w) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ void sts_pack2_x(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1, const float (&v)[4], float v4) {
    float hi[5], lo[5];
    split5(v, v4, hi, lo);
    sts4(a0, hi[0], hi[1], hi[2], hi[3]); sts4(a1, hi[4], lo[0], lo[1], lo[2]);
    sts4(b0, hi[0], hi[1], hi[2], hi[3]); sts4(b1, hi[4], lo[3], lo[4], 0.f);
}
__device__ __forceinline__ void sts_pack2_w(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1, const float (&v)[4], float v4) {
    float hi[5], lo[5];
    split5(v, v4, hi, lo);
    sts4(a0, hi[0], hi[1], hi[2], hi[3]); sts4(a1, hi[4], hi[0], hi[1], hi[2]);
    sts4(b0, lo[0], lo[1], lo[2], lo[3]); sts4(b1, lo[4], hi[3], hi[4], 0.f);
}

// kCompact: 0 = every centroid; 1 = live-person list (TcArgs::live_groups); 2 = live-TILE list (TcArgs::live_tiles);
//           3 = live-UNIT list (TcArgs::live_units): a tile is 16 octets of sample slots, all of one person
template <int C0, bool kTrain = false, int kCompact = 0>
__global__ void __launch_bounds__(kThreads, 1)
sa_mlp_tc16_kernel(const __grid_constant__ TcLayer1 L1, const __grid_constant__ TcBound BD, const __grid_constant__ TcArgs A) {
    static_assert(C0 >= 3 && C0 <= 7, "layer-1 K is padded to 8 with one column for the folded bias");
    // <= 5 used K columns (C0 <= 4): the 15 products xh.wh + xh.wl + xl.wh of the 3xTF32 split fit the 16 K-slots of
    // TWO MMAs instead of three: tile A = [xh0..xh4, xl0, xl1, xl2] . [wh0..wh4, wh0, wh1, wh2],
    //                            tile B = [xh0..xh4, xl3, xl4, 0  ] . [wl0..wl4, wh3, wh4, 0  ]
    constexpr bool kPack2 = C0 + 1 <= 5;
    extern __shared__ uint8_t smem_raw[];
    // full[4] empty[4] | tfull[2] tempty[2] | xfull[G] xempty[G] | dfull[G] dempty[G]
    __shared__ __align__(8) uint64_t s_bar[2 * kStages + 4 + 4 * kGroups];
    __shared__ __align__(16) float s_inv[kRing][4];     // 2^e of the four centroids of tile (i mod kRing)
    __shared__ uint32_t s_tmem;

    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B atoms need 1024 B alignment
    const uint32_t b_stage0 = base;                                 // h1 stage s: hi at +s*2*kHBytes, lo at +kHBytes more
    const uint32_t x_stage0 = base + kStages * 2 * kHBytes;         // x stage: hi at +xs*2*kXBytes, lo at +kXBytes more
    const uint32_t w1_hi = x_stage0 + kGroups * 2 * kXBytes, w1_lo = w1_hi + kW1Bytes;
    const uint32_t bar0 = smem_u32(s_bar);
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (kStages + s); };
    auto TFULL = [&](int t) { return bar0 + 8u * (2 * kStages + t); };
    auto TEMPTY = [&](int t) { return bar0 + 8u * (2 * kStages + 2 + t); };
    auto XFULL = [&](int x) { return bar0 + 8u * (2 * kStages + 4 + x); };
    auto XEMPTY = [&](int x) { return bar0 + 8u * (2 * kStages + 4 + kGroups + x); };
    auto DFULL = [&](int x) { return bar0 + 8u * (2 * kStages + 4 + 2 * kGroups + x); };
    auto DEMPTY = [&](int x) { return bar0 + 8u * (2 * kStages + 4 + 3 * kGroups + x); };

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // M counts "virtual" centroids when the live-person list is in use (TcArgs::live_groups): only they are walked
    constexpr bool compact = kCompact == 1;
    constexpr bool units = kCompact == 3;      // tile i of the walk = the 16 (centroid, octet) entries live_units[16 i ..] of ONE person
    constexpr bool tiled = kCompact == 2 || units;   // tile i = the four real centroids live_tiles[i] .. +3 of ONE person
    const int M = tiled ? __ldg(A.n_live) * 4 : (compact ? __ldg(A.n_live) * A.group_size : A.M);
    const int n_tiles = (M + 3) >> 2;
    // contiguous tile range per CTA: consecutive tiles belong to the same person, so the running
    // group max of the epilogue is flushed only a handful of times
    const int t_beg = (int)((long long)blockIdx.x * n_tiles / gridDim.x);
    const int t_end = (int)((long long)(blockIdx.x + 1) * n_tiles / gridDim.x);
    const int my_tiles = t_end - t_beg;

    // ---- one-time setup: barriers, TMEM, weight operands ---------------------------------------------
    if (tid == 0) {
        for (int s = 0; s < kStages; s++) { mbar_init(FULL(s), 128); mbar_init(EMPTY(s), 1); }
        for (int t = 0; t < kAcc; t++) { mbar_init(TFULL(t), 1); mbar_init(TEMPTY(t), 128); }
        for (int x = 0; x < kGroups; x++) {
            mbar_init(XFULL(x), 128); mbar_init(XEMPTY(x), 1);
            mbar_init(DFULL(x), 1); mbar_init(DEMPTY(x), 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // W1^T (64 channels x K=8: C0 folded weights, the folded BatchNorm shift in column C0, zeros) as the
    // B operand of the layer-1 MMA, tf32 hi/lo split, no-swizzle K-major
    if (tid >= 160 && tid < 224) {                     // (two converter warps)
        const int ch = tid - 160;
        float k0[4], k1[4];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            // (training statistics pass: BatchNorm 1's batch scale / shift only exist on the device - TcArgs::l1_dev)
            const float v = (kTrain && A.l1_dev)
                                ? (k < C0 ? __ldg(A.l1_dev + k * kTcC1 + ch) : (k == C0 ? __ldg(A.l1_dev + 8 * kTcC1 + ch) : 0.f))
                                : (k < C0 ? L1.w[k][ch] : (k == C0 ? L1.sh[ch] : 0.f));
            if (k < 4) k0[k] = v; else k1[k - 4] = v;
        }
        if (kPack2) sts_pack2_w(w1_hi + k8_off(ch, 0), w1_hi + k8_off(ch, 1), w1_lo + k8_off(ch, 0), w1_lo + k8_off(ch, 1), k0, k1[0]);
        else {
            sts_split4(w1_hi + k8_off(ch, 0), w1_lo + k8_off(ch, 0), k0);
            sts_split4(w1_hi + k8_off(ch, 1), w1_lo + k8_off(ch, 1), k1);
        }
        fence_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    // W2 lives in TENSOR MEMORY as the A operand of the layer-2 MMA (lane = output channel; a 32-bit column
    // holds the fp16 pair k = 2j, 2j+1), scaled into the fp16 range and sign-aligned with scale2 (see header).
    float asc = 0.f, shift = 0.f;        // epilogue: |scale2| * 2^-ec and shift2 of this thread's channel
    float sgn = 1.f;
    if (warp < 4) {
        const int c = tid;
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        const float sc = __ldg(A.sc2 + c);
        shift = __ldg(A.sh2 + c);
        float w[kTcC1], amax = 0.f;
#pragma unroll
        for (int k = 0; k < kTcC1; k++) { w[k] = __ldg(A.w2t + (size_t)k * kTcC2 + c); amax = fmaxf(amax, fabsf(w[k])); }
        int eb = (int)(__float_as_uint(amax) >> 23);
        eb = eb < 14 ? 14 : eb;                                               // (all-zero / denormal rows: any scale works)
        const float up = __uint_as_float((uint32_t)(267 - eb) << 23);        // 2^(13 - floor(log2 amax))
        const float down = __uint_as_float((uint32_t)(eb - 13) << 23);       // its inverse
        const float sup = sc < 0.f ? -up : up;
        sgn = sc < 0.f ? -1.f : 1.f;
        asc = __fmul_rn(fabsf(sc), down);
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int j = 0; j < 32; j++) split_h2(__fmul_rn(w[2 * j], sup), __fmul_rn(w[2 * j + 1], sup), hi[j], lo[j]);
        tc_st32(lane_base + kTmW2Hi, hi);
        tc_st32(lane_base + kTmW2Lo, lo);
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (warp < 4) {
        // ===== epilogue: thread = output channel, 32 consecutive TMEM columns = one centroid ==========
        const int c = tid;
        if constexpr (kTrain) {
            // training statistics (sa_mlp_train_fast.cu, identity 1): no BatchNorm yet - its statistics are what
            // is being gathered.  The accumulator holds sign(gamma2) * z2 * 2^(ec - e): extreme and its first
            // sample per centroid, sum and sum of squares per channel (fp32 over 32 samples, fp64 across).
            double dsum = 0.0, dsq = 0.0;
            int m_next = t_beg * 4;
            for (int i = 0; i < my_tiles; i++) {
                const int t = i % kAcc, u = i / kAcc;
                mbar_wait(TFULL(t), u & 1);
                tc_fence_after();
                const float4 inv = *reinterpret_cast<const float4 *>(&s_inv[i % kRing][0]);
                const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(t * 128);
                float zx[4], ps[4], pq[4];
                int ar[4];
                bool nr[4];
                {
                    uint32_t v0[32], v1[32];
                    tc_ld32(taddr, v0);
                    tc_ld32(taddr + 32, v1);
                    tc_wait_ld();
                    tr_reduce32(v0, zx[0], ar[0], ps[0], pq[0], nr[0]);
                    tr_reduce32(v1, zx[1], ar[1], ps[1], pq[1], nr[1]);
                    const uint32_t dep = (__float_as_uint(zx[0]) & __float_as_uint(zx[1])) == 0xffffffffu ? 1u : 0u;
                    tc_ld32(taddr + 64 + dep, v0);
                    tc_ld32(taddr + 96 + dep, v1);
                    tc_wait_ld();
                    tr_reduce32(v0, zx[2], ar[2], ps[2], pq[2], nr[2]);
                    tr_reduce32(v1, zx[3], ar[3], ps[3], pq[3], nr[3]);
                }
                tc_fence_before();
                mbar_arrive_relaxed(TEMPTY(t));
                const float sq[4] = {inv.x, inv.y, inv.z, inv.w};
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int m = m_next + q;
                    if (m < M) {
                        const float scale = __fmul_rn(asc, sq[q]);            // exact power of two
                        A.tr_zext[(size_t)m * kTcC2 + c] = __fmul_rn(zx[q], scale);
                        A.tr_arg[(size_t)m * kTcC2 + c] = ar[q] | (nr[q] ? 0x100 : 0);   // 0x100: near tie, see tr_reduce32
                        dsum += (double)ps[q] * (double)scale;
                        dsq += (double)pq[q] * (double)scale * (double)scale;
                    }
                }
                m_next += 4;
            }
            if (my_tiles > 0) {
                atomicAdd(A.tr_sum + c, (double)sgn * dsum);
                atomicAdd(A.tr_sq + c, dsq);
            }
        } else {
        const int gs = A.group_size;
        int m_next = t_beg * 4;                                   // first centroid of the next tile
        int grp = tiled ? -1 : (A.gmax ? m_next / gs : 0), g_end = (grp + 1) * gs;
        float gm = 0.f;
        float *outp = (A.out && !tiled) ? A.out + (size_t)m_next * A.ld_out + c : nullptr;
        auto flush = [&]() {
            const int rg = compact ? __ldg(A.live_groups + grp) : grp;     // virtual -> real person
            if (tiled || !(A.group_zero && __ldg(A.group_zero + rg)))       // (the tile list holds live persons only)
                atomicMax(reinterpret_cast<int *>(A.gmax + (size_t)rg * kTcC2 + c), __float_as_int(gm));
        };
        for (int i = 0; i < my_tiles; i++) {
            const int t = i % kAcc, u = i / kAcc;
            int tile_person = 0;
            if (tiled) {
                const int c0 = units ? (__ldg(A.live_units + (size_t)(t_beg + i) * kUnitsPerTile) & 0x7fffffff) >> kUnitShift
                                     : __ldg(A.live_tiles + t_beg + i);
                tile_person = A.m_shift >= 0 ? c0 >> A.m_shift : c0 / gs;
            }
            mbar_wait(TFULL(t), u & 1);
            if (warp == 0) TR(12, i);
            tc_fence_after();
            const float4 inv = *reinterpret_cast<const float4 *>(&s_inv[i % kRing][0]);
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(t * 128);
            float mx[4];
            {   // two centroids at a time (64 registers).  The second pair of loads is made to depend on the first
                // pair's maxima: hoisted above them (as ptxas does when left alone) it needs the same 64 registers
                // and ~45 MOVs per tile to shuffle the first pair out of the way; the epilogue has the slack.
                uint32_t v0[32], v1[32];
                tc_ld32(taddr, v0);
                tc_ld32(taddr + 32, v1);
                tc_wait_ld();
                mx[0] = max32(v0);
                mx[1] = max32(v1);
                // (a data dependence ptxas cannot fold: 0 unless both maxima are the all-ones NaN)
                const uint32_t dep = (__float_as_uint(mx[0]) & __float_as_uint(mx[1])) == 0xffffffffu ? 1u : 0u;
                tc_ld32(taddr + 64 + dep, v0);
                tc_ld32(taddr + 96 + dep, v1);
                tc_wait_ld();
                mx[2] = max32(v0);
                mx[3] = max32(v1);
            }
            tc_fence_before();
            mbar_arrive_relaxed(TEMPTY(t));
            if (warp == 0) TR(13, i);
            const float sq[4] = {inv.x, inv.y, inv.z, inv.w};
            // BatchNorm affine + ReLU of the max over the 32 samples (exact power-of-two rescale)
            float res[4];
#pragma unroll
            for (int q = 0; q < 4; q++) res[q] = fmaxf(__fmaf_rn(mx[q], __fmul_rn(asc, sq[q]), shift), 0.f);
            if (tiled) {
                if (tile_person != grp) {                     // the walk entered the next person
                    if (grp >= 0) flush();
                    grp = tile_person; gm = 0.f;
                }
                gm = fmaxf(max3(gm, res[0], res[1]), fmaxf(res[2], res[3]));
            } else if (m_next + 4 <= M && (!A.gmax || m_next + 4 <= g_end)) {
                // common case: four real centroids of one group
                if (outp) {
#pragma unroll
                    for (int q = 0; q < 4; q++) outp[(size_t)q * A.ld_out] = res[q];
                }
                gm = fmaxf(max3(gm, res[0], res[1]), fmaxf(res[2], res[3]));
            } else {
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
            }
            m_next += 4;
            if (outp) outp += (size_t)4 * A.ld_out;
        }
        if (A.gmax && my_tiles > 0) flush();
        }
    } else if (warp >= kWarpL2 && warp < kWarpL1) {
        // ===== layer-2 MMA issuers: warp 4 takes the even tiles, warp 14 the odd ones.  Per tile 12 x (128x128x16):
        // small terms first (W2lo*h1hi, W2hi*h1lo), then W2hi*h1hi.  K step kk: 8 TMEM columns of W2, 32 bytes
        // inside the 128B-swizzled row of h1.  The whole warp runs the warp-uniform control flow; only the
        // tcgen05 instructions are predicated on one elected lane.
        const uint32_t idesc2 = (1u << 4) /* D fp32 */ | (0u << 7) /* A fp16 */ | (0u << 10) /* B fp16 */ |
                                ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        for (int i = warp - kWarpL2; i < my_tiles; i += kIssuers) {
            const int s = i % kStages, j = i / kStages, t = i % kAcc, u = i / kAcc;
            mbar_wait(FULL(s), j & 1);
            TR(10, i);
            mbar_wait(TEMPTY(t), (u & 1) ^ 1);
            TR(11, i);
            tc_fence_after();
            const uint32_t b_hi = b_stage0 + (uint32_t)s * 2 * kHBytes;
            const uint64_t bd_hi = make_desc(b_hi), bd_lo = make_desc(b_hi + kHBytes);
            const uint32_t d = tmem + (uint32_t)(t * 128);
            if (elect_one()) {
#pragma unroll
                for (int term = (MGAR_TC16_EXP == 2 ? 1 : 0); term < 3; term++) {
                    const uint32_t a0 = tmem + (term == 0 ? kTmW2Lo : kTmW2Hi);
                    const uint64_t b0 = term == 1 ? bd_lo : bd_hi;
#pragma unroll
                    for (int kk = 0; kk < 4; kk++)
                        tc_mma_f16_ts(d, a0 + (uint32_t)(kk * 8), b0 + (uint64_t)((kk * 32) >> 4), idesc2,
                                      ((term - (MGAR_TC16_EXP == 2 ? 1 : 0)) | kk) ? 1u : 0u);
                }
                tc_commit(EMPTY(s));     // smem stage reusable once these MMAs have read it
                tc_commit(TFULL(t));     // accumulator ready for the epilogue
            }
            __syncwarp();
        }
    } else if (warp == kWarpL1) {
        // ===== layer-1 MMA issuer (3xTF32, K = 8): D1[pos, ch] = x[pos, 0:8] . W1^T[ch, 0:8] ===================
        const uint32_t idesc1 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint64_t w1d_hi = make_desc_k8(w1_hi), w1d_lo = make_desc_k8(w1_lo);
        for (int i = 0; i < my_tiles; i++) {
            const int xs = i % kGroups, jx = i / kGroups;
            mbar_wait(XFULL(xs), jx & 1);
            TR(8, i);
            mbar_wait(DEMPTY(xs), (jx & 1) ^ 1);
            TR(9, i);
            tc_fence_after();
            const uint32_t x_hi = x_stage0 + (uint32_t)xs * 2 * kXBytes;
            const uint64_t xd_hi = make_desc_k8(x_hi), xd_lo = make_desc_k8(x_hi + kXBytes);
            const uint32_t d1 = tmem + kTmD1 + (uint32_t)(xs * 64);
            if (elect_one()) {
                if (kXTmem) {                                   // x tile in TMEM: columns 0-7 = "hi" tile, 8-15 = "lo" tile
                    const uint32_t xa = tmem + kTmX + (uint32_t)(xs * 16);
                    if (kPack2) {
                        tc_mma_tf32_ts(d1, xa + 8, w1d_lo, idesc1, 0);
                        tc_mma_tf32_ts(d1, xa, w1d_hi, idesc1, 1);
                    } else {
                        tc_mma_tf32_ts(d1, xa + 8, w1d_hi, idesc1, 0);
                        tc_mma_tf32_ts(d1, xa, w1d_lo, idesc1, 1);
                        tc_mma_tf32_ts(d1, xa, w1d_hi, idesc1, 1);
                    }
                } else if (kPack2) {                            // 15 products in 2 x 8 K-slots (see sts_pack2_x)
                    tc_mma_tf32(d1, xd_lo, w1d_lo, idesc1, 0);
                    if (MGAR_TC16_EXP != 4) tc_mma_tf32(d1, xd_hi, w1d_hi, idesc1, 1);
                } else {
                    tc_mma_tf32(d1, xd_lo, w1d_hi, idesc1, 0);
                    tc_mma_tf32(d1, xd_hi, w1d_lo, idesc1, 1);
                    tc_mma_tf32(d1, xd_hi, w1d_hi, idesc1, 1);
                }
                tc_commit(XEMPTY(xs));
                tc_commit(DFULL(xs));
            }
            __syncwarp();
        }
    } else {
        // ===== gather + converters: two groups of four warps, group g owns the tiles i = g (mod 2) ===========
        // warp = (group g, position block pw = one centroid = one TMEM lane quarter).  A group has two tile
        // periods per tile of its own, so the serial chain of a tile (TMEM load -> convert -> swizzled store)
        // is off the critical path of the tensor pipe.  Per own tile i:
        //   (a) gather (xyz - centre | feat) of tile i+2, derive the centroid's power-of-two scale, split hi/lo
        //       (tf32) and store the 128 x 8 layer-1 input tile (bias column = the scale) into x stage g; the
        //       gather itself is software-pipelined two more own tiles ahead in registers;
        //   (b) read the position's 64 layer-1 outputs back from TMEM (D1 stage g), ReLU, split into fp16
        //       hi/lo and store the 128B-swizzled K-major h1 operand of the layer-2 MMA.
        const int pw = warp & 3, g = (warp - 4) >> 2;
        const int r = pw * 32 + lane;

        int u_cloud = A.uniform ? (t_beg * 4 + pw) / A.m_u : 0, u_end = (u_cloud + 1) * A.m_u;
        int i_cloud = u_cloud, i_end = u_end;                    // the same walk for the index loads (they run ahead)
        // real person of the current virtual one, re-read only when the walk enters the next person (every 128 tiles)
        // kept as offsets so a tile pays one add: real centroid = virtual + delta, cloud start = u_start
        int u_delta = 0, i_delta = 0;
        long long u_start = 0;
        if (compact && my_tiles > 0) {
            const int real = __ldg(A.live_groups + u_cloud);
            u_delta = i_delta = (real - u_cloud) * A.m_u;
            u_start = (long long)real * A.n_u;
        }
        // ns != 32: the walk's unit m is a BLOCK of 32 positions (TcArgs::ns); cen(m) = the centroid this lane's position
        // belongs to (ns = 16: per half warp)
        const int ns = A.ns;
        auto cen = [&](int m) { return ns == kTcNS ? m : (ns == 64 ? (m >> 1) : 2 * m + (lane >> 4)); };
        // bound of |layer-1 pre-activation| (TcBound): kernel arguments, or - training - the nine floats behind l1_dev's W / shift
        float bd_w[C0], bd_sh;
        {
            const bool dev = kTrain && A.l1_dev;
#pragma unroll
            for (int cc = 0; cc < C0; cc++) bd_w[cc] = dev ? __ldg(A.l1_dev + 9 * kTcC1 + cc) : BD.wmax[cc];
            bd_sh = dev ? __ldg(A.l1_dev + 9 * kTcC1 + 8) : BD.shmax;
        }
        int ent_next = 0;                                        // live-unit walk: this lane's list entry of the next load_idx call
        if (units && g < my_tiles) ent_next = __ldg(A.live_units + (size_t)(t_beg + g) * kUnitsPerTile + pw * (kUnitsPerTile / 4) + lane / kUnitSlots);
        auto load_idx = [&](int i, int &iv, int &em, int &mr) {   // local tile i -> this lane's sample index (mr: real centroid)
            int m = (t_beg + i) * 4 + pw;
            iv = -1; em = 0; mr = 0;
            if (units) {                                      // lane -> octet pw * 4 + lane / 8 of the tile, slot lane % 8 of it
                // the entry was fetched by the previous call (the calls walk this group's tiles in order, i += kGroups): a tile's
                // sixteen entries are a 64-byte line of their own, i.e. a trip to L2 that must not sit in front of the index load
                const int e = ent_next;
                if (i + kGroups < my_tiles)
                    ent_next = __ldg(A.live_units + (size_t)(t_beg + i + kGroups) * kUnitsPerTile + pw * (kUnitsPerTile / 4) + lane / kUnitSlots);
                if (i < my_tiles) {
                    mr = (e & 0x7fffffff) >> kUnitShift;
                    em = e < 0;                               // bit 31: empty ball (the builder saw -1 in slot 0)
                    iv = __ldg(A.idx + (size_t)mr * kTcNS + (e & (kUnitsPerRow - 1)) * kUnitSlots + (lane & (kUnitSlots - 1)));
                }
                return;
            }
            if (tiled) {
                if (i < my_tiles) {
                    mr = __ldg(A.live_tiles + t_beg + i) + pw;
                    iv = __ldg(A.idx + (size_t)mr * kTcNS + lane);
                    if (A.empty_mask) em = A.empty_mask[cen(mr)];
                    if (ns == 64 && (mr & 1) && __ldg(A.idx + (size_t)(mr >> 1) * 64) < 0) em = 1;
                }
                return;
            }
            if (i < my_tiles && m < M) {
                if (compact) {                                   // virtual -> real centroid (m_u == group_size)
                    while (m >= i_end) {
                        i_cloud++; i_end += A.m_u;
                        i_delta = (__ldg(A.live_groups + i_cloud) - i_cloud) * A.m_u;
                    }
                    m += i_delta;
                }
                iv = __ldg(A.idx + (size_t)m * kTcNS + lane);
                if (A.empty_mask) em = A.empty_mask[cen(m)];
                if (ns == 64 && (m & 1) && __ldg(A.idx + (size_t)(m >> 1) * 64) < 0) em = 1;
            }
        };
        auto load_pt = [&](int i, int iv, int em, int mr, float (&v)[C0], float (&q)[3], bool &live) {
            int m = tiled ? mr : (t_beg + i) * 4 + pw;
            // slot 0 of the centroid's row (-1: empty ball): lane 0 of a whole centroid, lane 0 / 16 of a pair of
            // 16-sample centroids; the second block of a 64-sample centroid got it as `em` from load_idx
            const int first = units ? 0 : __shfl_sync(0xffffffffu, iv, ns == 16 ? (lane & 16) : 0);
            live = (i < my_tiles) && (tiled || m < M) && first >= 0 && !em;
            if (live) {
                long long start;
                if (tiled) {
                    start = (long long)(A.m_shift >= 0 ? m >> A.m_shift : m / A.m_u) * A.n_u;
                } else if (A.uniform) {
                    if (compact) {
                        while (m >= u_end) {
                            u_cloud++; u_end += A.m_u;
                            const int real = __ldg(A.live_groups + u_cloud);
                            u_delta = (real - u_cloud) * A.m_u;
                            u_start = (long long)real * A.n_u;
                        }
                        m += u_delta;
                        start = u_start;
                    } else {
                        while (m >= u_end) { u_cloud++; u_end += A.m_u; }     // tiles arrive in increasing order
                        start = (long long)u_cloud * A.n_u;
                    }
                } else {
                    start = __ldg(A.pre_xyz + find_batch_tc(A.pre_new, A.B, cen(m)));
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
                const size_t mc = (size_t)cen(m) * 3;
                q[0] = __ldg(A.new_xyz + mc + 0);
                q[1] = __ldg(A.new_xyz + mc + 1);
                q[2] = __ldg(A.new_xyz + mc + 2);
            }
        };
        // layer-1 input row of local tile i -> x stage i & 1, scaled by the centroid's 2^-e
        // publish = false (kDeferX): the caller makes the tile visible later (publish_x), once the stores have long completed
        auto publish_x = [&](int i) {
            if (!kXTmem) fence_async_smem();
            mbar_arrive(XFULL(i % kGroups));
        };
        auto store_x = [&](int i, const float (&v)[C0], const float (&q)[3], bool live, bool publish) {
            const int xs = i % kGroups, jx = i / kGroups;    // xs == g for this group's tiles
            float x[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (live) {
                x[0] = __fsub_rn(v[0], q[0]); x[1] = __fsub_rn(v[1], q[1]); x[2] = __fsub_rn(v[2], q[2]);
#pragma unroll
                for (int cc = 3; cc < C0; cc++) x[cc] = v[cc];
            }
            float bound = bd_sh;
#pragma unroll
            for (int cc = 0; cc < C0; cc++) bound = __fmaf_rn(bd_w[cc], fabsf(x[cc]), bound);
            // bound >= 0: the bit patterns order like the values.  bound < 2^(eb - 126); 2^(141 - eb) puts it below 2^15.
            int eb = (int)(__reduce_max_sync(0xffffffffu, __float_as_uint(bound)) >> 23);
            eb = eb < 15 ? 15 : (eb > 254 ? 254 : eb);
            const float down = __uint_as_float((uint32_t)(268 - eb) << 23);    // 2^(141 - eb)
            const float up = __uint_as_float((uint32_t)(eb - 14) << 23);       // 2^(eb - 141)
#pragma unroll
            for (int cc = 0; cc < C0; cc++) x[cc] = __fmul_rn(x[cc], down);
            x[C0] = down;                                        // multiplies the folded BatchNorm shift
            const float k0[4] = {x[0], x[1], x[2], x[3]}, k1[4] = {x[4], x[5], x[6], x[7]};
            const uint32_t x_hi = x_stage0 + (uint32_t)xs * 2 * kXBytes, x_lo = x_hi + kXBytes;
            if (pw == 0) TR(0, i);
            mbar_wait(XEMPTY(xs), (jx & 1) ^ 1);
            if (pw == 0) TR(1, i);
            if (kXTmem) {
                tc_fence_after();                                // the MMA that read this stage has completed (XEMPTY)
                uint32_t xv[16];
                if (kPack2) {                                    // the two packed tiles of sts_pack2_x, one row each
                    float hi[5], lo[5];
                    split5(k0, k1[0], hi, lo);
                    const float ta[8] = {hi[0], hi[1], hi[2], hi[3], hi[4], lo[0], lo[1], lo[2]};
                    const float tb[8] = {hi[0], hi[1], hi[2], hi[3], hi[4], lo[3], lo[4], 0.f};
#pragma unroll
                    for (int q = 0; q < 8; q++) { xv[q] = __float_as_uint(ta[q]); xv[8 + q] = __float_as_uint(tb[q]); }
                } else {                                         // tf32 truncation / exact remainder, as sts_split4
#pragma unroll
                    for (int q = 0; q < 8; q++) {
                        const float hq = __uint_as_float(__float_as_uint(x[q]) & 0xffffe000u);
                        xv[q] = __float_as_uint(hq);
                        xv[8 + q] = __float_as_uint(__fsub_rn(x[q], hq));
                    }
                }
                tc_st16(tmem + ((uint32_t)(pw * 32) << 16) + kTmX + (uint32_t)(xs * 16), xv);
                tc_wait_st();
                tc_fence_before();
            } else if (MGAR_TC16_EXP == 3) asm volatile("" ::"f"(k0[0] + k0[1] + k0[2] + k0[3] + k1[0]));
            else if (kPack2) sts_pack2_x(x_hi + k8_off(r, 0), x_hi + k8_off(r, 1), x_lo + k8_off(r, 0), x_lo + k8_off(r, 1), k0, k1[0]);
            else {
                sts_split4(x_hi + k8_off(r, 0), x_lo + k8_off(r, 0), k0);
                sts_split4(x_hi + k8_off(r, 1), x_lo + k8_off(r, 1), k1);
            }
            if (lane == 0) s_inv[i % kRing][pw] = up;
            if (publish) publish_x(i);
            if (pw == 0) TR(2, i);
        };

        int iv_a = -1, em_a = 0, iv_b = -1, em_b = 0, mr_a = 0, mr_b = 0;
        float vb[C0], qb[3];
        bool live_b = false;
        load_idx(g, iv_b, em_b, mr_b);
        load_pt(g, iv_b, em_b, mr_b, vb, qb, live_b);
        load_idx(g + kGroups, iv_b, em_b, mr_b);
        load_idx(g + 2 * kGroups, iv_a, em_a, mr_a);
        if (g < my_tiles) store_x(g, vb, qb, live_b, true);
        load_pt(g + kGroups, iv_b, em_b, mr_b, vb, qb, live_b);
        for (int i = g, k = 0; i < my_tiles; i += kGroups, k++) {
            if (i + kGroups < my_tiles) {
                store_x(i + kGroups, vb, qb, live_b, !kDeferX); // gathered during the previous iteration
                iv_b = iv_a; em_b = em_a; mr_b = mr_a;
                load_pt(i + 2 * kGroups, iv_b, em_b, mr_b, vb, qb, live_b);   // points two own tiles ahead
                load_idx(i + 3 * kGroups, iv_a, em_a, mr_a);                  // indices three own tiles ahead
            }
            const int s = i % kStages, j = i / kStages;
            if (pw == 0) TR(3, i);
            mbar_wait(DFULL(g), k & 1);
            if (pw == 0) TR(4, i);
            tc_fence_after();
            uint32_t h0[32], h1[32];
            const uint32_t d1 = tmem + ((uint32_t)(pw * 32) << 16) + kTmD1 + (uint32_t)(g * 64);
            tc_ld32(d1, h0);
            tc_ld32(d1 + 32, h1);
            tc_wait_ld();
            tc_fence_before();
            mbar_arrive_relaxed(DEMPTY(g));                     // layer 1 of tile i+2 may overwrite D1 now
            if (kDeferX && i + kGroups < my_tiles) publish_x(i + kGroups);   // ... and its input tile, stored at the top of this iteration
            if (pw == 0) TR(5, i);

            const uint32_t b_hi = b_stage0 + (uint32_t)s * 2 * kHBytes, b_lo = b_hi + kHBytes;
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int p = 0; p < 16; p++)
                split_relu_h2(__uint_as_float(h0[2 * p]), __uint_as_float(h0[2 * p + 1]), hi[p], lo[p]);
            mbar_wait(EMPTY(s), (j & 1) ^ 1);
            if (pw == 0) TR(6, i);
#pragma unroll
            for (int half = 0; half < 2; half++) {
                if (half == 1) {
#pragma unroll
                    for (int p = 0; p < 16; p++)
                        split_relu_h2(__uint_as_float(h1[2 * p]), __uint_as_float(h1[2 * p + 1]), hi[p], lo[p]);
                }
#pragma unroll
                for (int c4 = 0; c4 < 4; c4++) {
                    const uint32_t off = h_off(r, half * 4 + c4);
                    if (MGAR_TC16_EXP == 1) { asm volatile("" ::"r"(hi[4 * c4] ^ lo[4 * c4 + 1] ^ hi[4 * c4 + 2] ^ lo[4 * c4 + 3] ^ hi[4 * c4 + 1] ^ lo[4 * c4] ^ hi[4 * c4 + 3] ^ lo[4 * c4 + 2])); continue; }
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(b_hi + off), "r"(hi[4 * c4]), "r"(hi[4 * c4 + 1]),
                                 "r"(hi[4 * c4 + 2]), "r"(hi[4 * c4 + 3]) : "memory");
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(b_lo + off), "r"(lo[4 * c4]), "r"(lo[4 * c4 + 1]),
                                 "r"(lo[4 * c4 + 2]), "r"(lo[4 * c4 + 3]) : "memory");
                }
            }
            fence_async_smem();          // generic-proxy stores -> visible to the tensor core (async proxy)
            mbar_arrive(FULL(s));
            if (pw == 0) TR(7, i);
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

}  // namespace

// live[0 .. *n_live) = ascending g with group_zero[g] == 0: one CTA, 1024 groups per pass (ballot + warp totals)
__global__ void __launch_bounds__(1024)
compact_live_groups_kernel(const int *__restrict__ group_zero, int G, int *__restrict__ live, int *__restrict__ n_live) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int g0 = 0; g0 < G; g0 += 1024) {
        const int g = g0 + tid;
        const bool keep = g < G && __ldg(group_zero + g) == 0;
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int before = 0;
        for (int w = 0; w < warp; w++) before += s_warp[w];
        if (keep) live[s_base + before + __popc(bal & ((1u << lane) - 1u))] = g;
        __syncthreads();
        if (tid == 1023) s_base += before + __popc(bal);
        __syncthreads();
    }
    if (tid == 0) *n_live = s_base;
}

int tc_compact_live_groups(const int *group_zero, int G, int *live, int *n_live, cudaStream_t st) {
    compact_live_groups_kernel<<<1, 1024, 0, st>>>(group_zero, G, live, n_live);
    count_launch();
    return check_launch("compact_live_groups");
}

// CTA c owns the persons [256 c, 256 c + 256): it sums the tile counts of all the persons before them (a coalesced
// read of <= G ints - cheaper than a second kernel or a look-back chain for a few thousand persons), scans its own 256
// counts and writes its tile rows, a warp per person, lanes = consecutive tiles.  (The first version - one CTA looping
// over all persons - took 27 us per step, 3 % of it.)
constexpr int kTilesCta = 256;
__device__ __forceinline__ int person_tiles(const int *__restrict__ group_zero, const int *__restrict__ live_cnt, int g, int gs, int ns) {
    if (__ldg(group_zero + g) != 0) return 0;
    int lc = (__ldg(live_cnt + g) * ns + 31) >> 5;                   // centroids -> blocks of 32 positions (gs counts blocks)
    lc = lc < 1 ? 1 : (lc > gs ? gs : lc);
    return (lc + 3) >> 2;
}
__global__ void __launch_bounds__(kTilesCta)
build_live_tiles_kernel(const int *__restrict__ group_zero, const int *__restrict__ live_cnt, int G, int gs, int ns,
                        int *__restrict__ tiles, int *__restrict__ n_tiles) {
    __shared__ int s_warp[kTilesCta / 32];
    __shared__ int s_first[kTilesCta], s_cnt[kTilesCta];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g0 = blockIdx.x * kTilesCta;
    int before = 0;
    for (int g = tid; g < g0; g += kTilesCta) before += person_tiles(group_zero, live_cnt, g, gs, ns);
    before = __reduce_add_sync(0xffffffffu, before);
    if (tid == 0) s_base = 0;
    __syncthreads();
    if (lane == 0 && before) atomicAdd(&s_base, before);
    const int g = g0 + tid;
    const int c = g < G ? person_tiles(group_zero, live_cnt, g, gs, ns) : 0;
    int incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int first = s_base + incl - c;
    for (int w = 0; w < warp; w++) first += s_warp[w];
    s_first[tid] = first;
    s_cnt[tid] = c;
    __syncthreads();
    for (int q = warp; q < kTilesCta && g0 + q < G; q += kTilesCta / 32) {
        const int f = s_first[q], cnt = s_cnt[q], c0 = (g0 + q) * gs;
        for (int k = lane; k < cnt; k += 32) tiles[f + k] = c0 + 4 * k;
    }
    if (g == G - 1) *n_tiles = first + c;
}

int tc_build_live_tiles(const int *group_zero, const int *live_cnt, int G, int group_size, int ns, int *tiles, int *n_tiles,
                        cudaStream_t st) {
    if (G <= 0) return MGAR_OK;
    build_live_tiles_kernel<<<(G + kTilesCta - 1) / kTilesCta, kTilesCta, 0, st>>>(group_zero, live_cnt, G, group_size, ns, tiles, n_tiles);
    count_launch();
    return check_launch("build_live_tiles");
}

// ---- live-UNIT list (TcArgs::live_units) ------------------------------------------------------------------------------
// One CTA per person g.  A ball row holds the first hits in index order, then - in a wrapped cloud of p distinct points
// - the COPIES of those hits (indices >= p), then the first hit again as padding (ball_query_gpu.cu:49-60): only the
// leading entries that are < p and differ from entry 0 are distinct samples, the others repeat one of them and cannot
// change a max.  So centroid k needs u_k = 1 + [slot 8 distinct] + [slot 16 distinct] + [slot 24 distinct] octets of
// its 32 slots (distinct entries are a prefix of the row).  The person's octets are packed 16 to a tile (padded with
// copies of its first octet), the tiles appended to the list with one atomicAdd per person: the order of the persons in
// the list is whatever the atomics give - it shapes the schedule, never a result (per-person maxima are atomicMax'ed).
constexpr int kUnitThreads = 128, kUnitCache = 1024;
// octets of centroid row r (32 slots) in a cloud of p distinct points; e0 = slot 0 (-1: empty ball -> one octet of zeros)
__device__ __forceinline__ int row_octets(const int *__restrict__ r, int p, int &e0) {
    e0 = __ldg(r);
    int e[kUnitsPerRow];
#pragma unroll
    for (int q = 1; q < kUnitsPerRow; q++) e[q] = __ldg(r + q * kUnitSlots);       // (all in flight together)
    int u = 1;
    bool on = e0 >= 0;
#pragma unroll
    for (int q = 1; q < kUnitsPerRow; q++) {
        on = on && e[q] < p && e[q] != e0;                  // distinct entries are a prefix of the row
        u += on ? 1 : 0;
    }
    return u;
}
__global__ void __launch_bounds__(kUnitThreads)
build_live_units_kernel(const int *__restrict__ group_zero, const int *__restrict__ live_cnt, const int *__restrict__ point_cnt,
                        const int *__restrict__ idx, int gs, int n_u, int *__restrict__ units, int *__restrict__ n_tiles) {
    __shared__ int s_warp[kUnitThreads / 32];
    __shared__ int s_base, s_total;
    __shared__ unsigned char s_u[kUnitCache];
    const int g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (__ldg(group_zero + g) != 0) return;
    int lc = __ldg(live_cnt + g);
    lc = lc < 1 ? 1 : (lc > gs ? gs : lc);
    int p = __ldg(point_cnt + g);
    p = p < 1 ? 1 : (p > n_u ? n_u : p);
    const int *rows = idx + (size_t)g * gs * kTcNS;
    // sweep 1: octets per centroid (kept in shared memory: bits 0-2 the count, bit 7 the empty-ball flag) and the
    // person's total -> its place in the list (one atomic per person)
    int mine = 0, e0;
    for (int k = tid; k < lc; k += kUnitThreads) {
        const int u = row_octets(rows + (size_t)k * kTcNS, p, e0);
        if (k < kUnitCache) s_u[k] = (unsigned char)(u | (e0 < 0 ? 0x80 : 0));
        mine += u;
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if (tid == 0) s_total = 0;
    __syncthreads();
    if (lane == 0) atomicAdd(&s_total, mine);
    __syncthreads();
    const int total = s_total;
    if (tid == 0) s_base = atomicAdd(n_tiles, (total + kUnitsPerTile - 1) / kUnitsPerTile) * kUnitsPerTile;
    __syncthreads();
    const int base = s_base;
    // sweep 2: the counts scanned, the entries written
    int running = 0;
    for (int k0 = 0; k0 < lc; k0 += kUnitThreads) {
        const int k = k0 + tid;
        int u = 0;
        bool dead = false;
        if (k < lc) {
            if (k < kUnitCache) { u = s_u[k] & 15; dead = (s_u[k] & 0x80) != 0; }
            else { u = row_octets(rows + (size_t)k * kTcNS, p, e0); dead = e0 < 0; }
        }
        int incl = u;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        __syncthreads();                                   // s_warp of the previous pass has been read
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        int pos = running + incl - u;
        for (int w = 0; w < warp; w++) pos += s_warp[w];
        for (int w = 0; w < kUnitThreads / 32; w++) running += s_warp[w];
        const int ent = (int)(((unsigned)(g * gs + k) << kUnitShift) | (dead ? 0x80000000u : 0u));
        for (int o = 0; o < u; o++) units[base + pos + o] = ent | o;
    }
    // padding of the last tile: copies of the person's first octet
    const int f0 = __ldg(rows);
    const int ent0 = (int)(((unsigned)(g * gs) << kUnitShift) | (f0 < 0 ? 0x80000000u : 0u));
    for (int q = total + tid; q < (total + kUnitsPerTile - 1) / kUnitsPerTile * kUnitsPerTile; q += kUnitThreads) units[base + q] = ent0;
}

int tc_units_per_row() { return kUnitsPerRow; }
int tc_units_per_tile() { return kUnitsPerTile; }

int tc_build_live_units(const int *group_zero, const int *live_cnt, const int *point_cnt, const int *idx, int G, int group_size,
                        int n_u, int *units, int *n_tiles, cudaStream_t st) {
    if (G <= 0) return MGAR_OK;
    MGAR_CUDA(cudaMemsetAsync(n_tiles, 0, sizeof(int), st));
    build_live_units_kernel<<<G, kUnitThreads, 0, st>>>(group_zero, live_cnt, point_cnt, idx, group_size, n_u, units, n_tiles);
    count_launch();
    return check_launch("build_live_units");
}

int sa_mlp_tc16_launch(const TcLayer1 &L1, const TcArgs &A, int C0, int grid, cudaStream_t st) {
    TcBound BD;
    BD.shmax = 0.f;
    for (int c = 0; c < 8; c++) BD.wmax[c] = 0.f;
    for (int o = 0; o < kTcC1; o++) {
        for (int c = 0; c < C0; c++) BD.wmax[c] = fmaxf(BD.wmax[c], fabsf(L1.w[c][o]));
        BD.shmax = fmaxf(BD.shmax, fabsf(L1.sh[o]));
    }
#ifdef MGAR_TC16_TRACE
#define TC16_TRACE_DUMP                                                                        \
    {                                                                                          \
        static long long h[32 * 16];                                                           \
        cudaStreamSynchronize(st);                                                             \
        cudaMemcpyFromSymbol(h, g_tc16_trace, sizeof(h));                                      \
        for (int i = 0; i < 32; i++) {                                                         \
            fprintf(stderr, "tc16trace %d", 100 + i);                                          \
            for (int q = 0; q < 14; q++) fprintf(stderr, " %lld", h[i * 16 + q] ? h[i * 16 + q] - h[10] : -1); \
            fprintf(stderr, "\n");                                                             \
        }                                                                                      \
    }
#else
#define TC16_TRACE_DUMP
#endif
    if (A.live_units && C0 == 4) {       // live-unit walk: only the distinct samples of the distinct centroids of the live persons
        MGAR_CUDA(cudaFuncSetAttribute(sa_mlp_tc16_kernel<4, false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        sa_mlp_tc16_kernel<4, false, 3><<<grid, kThreads, kSmem, st>>>(L1, BD, A);
        count_launch();
        TC16_TRACE_DUMP
        return check_launch("sa_mlp_tc16 (live units)");
    }
    if (A.live_tiles && C0 == 4) {       // live-tile walk: only the distinct centroids of the persons that pooled points
        MGAR_CUDA(cudaFuncSetAttribute(sa_mlp_tc16_kernel<4, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        sa_mlp_tc16_kernel<4, false, 2><<<grid, kThreads, kSmem, st>>>(L1, BD, A);
        count_launch();
        TC16_TRACE_DUMP
        return check_launch("sa_mlp_tc16 (live tiles)");
    }
    if (A.live_groups && C0 == 4) {      // live-person walk (the LiDAR branch: xyz + intensity)
        MGAR_CUDA(cudaFuncSetAttribute(sa_mlp_tc16_kernel<4, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        sa_mlp_tc16_kernel<4, false, 1><<<grid, kThreads, kSmem, st>>>(L1, BD, A);
        count_launch();
        return check_launch("sa_mlp_tc16 (live persons)");
    }
#define MGAR_TC16_CASE(c0)                                                                                          \
    if (C0 == c0) {                                                                                                 \
        MGAR_CUDA(cudaFuncSetAttribute(sa_mlp_tc16_kernel<c0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem)); \
        sa_mlp_tc16_kernel<c0><<<grid, kThreads, kSmem, st>>>(L1, BD, A);                                           \
        count_launch();                                                                                             \
        TC16_TRACE_DUMP                                                                                             \
        return check_launch("sa_mlp_tc16");                                                                         \
    }
    MGAR_TC16_CASE(3) MGAR_TC16_CASE(4) MGAR_TC16_CASE(5) MGAR_TC16_CASE(6) MGAR_TC16_CASE(7)
#undef MGAR_TC16_CASE
    set_error("sa_mlp_tc16: unsupported C0=%d", C0);
    return MGAR_EINVAL;
}

int sa_mlp_tc16_launch_train(const TcLayer1 &L1, const TcArgs &A, int C0, int grid, cudaStream_t st) {
    TcBound BD;
    BD.shmax = 0.f;
    for (int c = 0; c < 8; c++) BD.wmax[c] = 0.f;
    if (!A.l1_dev) {                     // (with l1_dev the kernel reads layer 1 and its bound from the device: L1 is unused)
        for (int o = 0; o < kTcC1; o++) {
            for (int c = 0; c < C0; c++) BD.wmax[c] = fmaxf(BD.wmax[c], fabsf(L1.w[c][o]));
            BD.shmax = fmaxf(BD.shmax, fabsf(L1.sh[o]));
        }
    }
#define MGAR_TC16_TRAIN_CASE(c0)                                                                                          \
    if (C0 == c0) {                                                                                                       \
        MGAR_CUDA(cudaFuncSetAttribute(sa_mlp_tc16_kernel<c0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem)); \
        sa_mlp_tc16_kernel<c0, true><<<grid, kThreads, kSmem, st>>>(L1, BD, A);                                           \
        count_launch();                                                                                                   \
        return check_launch("sa_mlp_tc16 (training statistics)");                                                         \
    }
    MGAR_TC16_TRAIN_CASE(4) MGAR_TC16_TRAIN_CASE(5)
#undef MGAR_TC16_TRAIN_CASE
    set_error("sa_mlp_tc16 (training statistics): unsupported C0=%d", C0);
    return MGAR_EINVAL;
}

}  // namespace mgar
