// fp32-faithful GEMM on tcgen05 (3xTF32):  C (M,N) = act(A (M,K) . Bt (N,K)^T + bias),  fp32 in, fp32 out.
//
// The dense contractions AROUND the LiDAR hot path that the reference leaves to cuBLAS / cuDNN in fp32:
//   * the ten 512 x 512 linears of every FusionAttention_mat module (model/gat_model.py:466-511: WQ/WK/WV of both
//     streams and the two FFN linears per stream) - batched here as (B*n, 512) x (512, 1536 | 512) products,
//   * the 128 -> 512 embedding of the LiDAR feature (gat_model.py:883-894),
//   * the [C_in -> C_out] 1x1 convolutions of NeighborVoxelSAModuleMSG (voxel_pool_modules.py:70-130) when their widths
//     are multiples of 32 / 128.
// torch runs them as SIMT sgemm (cutlass3x_sm100_simt_sgemm, profiles/r01_launches_config3.md: 21 % of the configs[2] step).
//
// Scheme (the one of sa_mlp_tc.cu, as a plain GEMM): every fp32 operand value is split v = hi + lo with hi = v truncated to
// tf32 (exact: lo is the 13 dropped mantissa bits) and the product is accumulated as lo.hi + hi.lo + hi.hi in the tensor
// core's fp32 accumulator; the dropped lo.lo term is 2^-22 relative.  kind::tf32, M = 128 (TMEM lanes = rows of A),
// N = 128 (columns), K = 8 per instruction, both operands K-major in shared memory with the 128-byte swizzle (a K chunk =
// 32 floats = one 128-byte row; hi and lo tiles of 16 KB each per operand), fp32 accumulator in 128 TMEM columns.
//
// CTA = one 128 x 128 output tile, 9 warps:  0-3 epilogue (tcgen05.ld, bias, ReLU, 512-byte row segments to global) |
// 4 TMEM allocation + MMA issue (one elected lane) | 5-8 producers: thread = one row of the A tile and one row of the Bt
// tile; it streams its 128-byte row segments of the next K chunk into registers (global -> register prefetch one chunk
// ahead), splits them and stores hi / lo into a 3-stage shared-memory ring (full / empty mbarriers; the empty barrier is
// armed by tcgen05.commit once the MMAs of the stage have read it).
#include <stdlib.h>

#include "mgar_common.cuh"
#include "tc_ptx.cuh"

namespace mgar {
namespace {

constexpr int kGmThreads = 9 * 32;
constexpr int kGmStages = 3;
constexpr int kGmKC = 32;                                 // floats per K chunk (one 128-byte swizzled row)
constexpr uint32_t kGmTile = 128 * kGmKC * 4;             // 16 KB: one 128 x 32 fp32 tile
constexpr uint32_t kGmStageBytes = 4 * kGmTile;           // A hi, A lo, B hi, B lo
constexpr uint32_t kGmSmem = kGmStages * kGmStageBytes + 1024;

__device__ __forceinline__ uint32_t gm_off(int r, int c16) {      // 16-byte chunk c16 of row r in a swizzled tile
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4));
}

__global__ void __launch_bounds__(kGmThreads, 1)
gemm_tf32x3_kernel(int M, int N, int K, const float *__restrict__ A, int lda, const float *__restrict__ Bt, int ldb,
                   const float *__restrict__ bias, int relu, float *__restrict__ C, int ldc) {
    extern __shared__ uint8_t gm_smem[];
    __shared__ __align__(8) uint64_t s_bar[2 * kGmStages + 1];
    __shared__ uint32_t s_tmem;
    const uint32_t base = (smem_u32(gm_smem) + 1023u) & ~1023u;   // SWIZZLE_128B atoms need 1024 B alignment
    const uint32_t bar0 = smem_u32(s_bar);
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (kGmStages + s); };
    const uint32_t TFULL = bar0 + 8u * (2 * kGmStages);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.x * 128, m0 = blockIdx.y * 128;
    const int n_chunks = K / kGmKC;

    if (tid == 0) {
        for (int s = 0; s < kGmStages; s++) { mbar_init(FULL(s), 128); mbar_init(EMPTY(s), 1); }
        mbar_init(TFULL, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (warp < 4) {
        // ===== epilogue: thread = output row (TMEM lane), 4 x 32 columns ======================================
        mbar_wait(TFULL, 0);
        tc_fence_after();
        const int m = m0 + warp * 32 + lane;
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
        for (int q = 0; q < 4; q++) {
            uint32_t v[32];
            tc_ld32(taddr + (uint32_t)(q * 32), v);
            tc_wait_ld();
            if (m < M) {
                float *out = C + (size_t)m * ldc + n0 + q * 32;
                const float *bq = bias ? bias + n0 + q * 32 : nullptr;
#pragma unroll
                for (int c = 0; c < 32; c += 4) {
                    float4 o = make_float4(__uint_as_float(v[c]), __uint_as_float(v[c + 1]), __uint_as_float(v[c + 2]),
                                           __uint_as_float(v[c + 3]));
                    if (bq) {
                        const float4 bb = __ldg(reinterpret_cast<const float4 *>(bq + c));
                        o.x = __fadd_rn(o.x, bb.x); o.y = __fadd_rn(o.y, bb.y); o.z = __fadd_rn(o.z, bb.z); o.w = __fadd_rn(o.w, bb.w);
                    }
                    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
                    *reinterpret_cast<float4 *>(out + c) = o;
                }
            }
        }
        tc_fence_before();
    } else if (warp == 4) {
        // ===== MMA issue: per K chunk 3 terms x 4 K-steps of 128 x 128 x 8 ===================================
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        for (int kc = 0; kc < n_chunks; kc++) {
            const int s = kc % kGmStages, j = kc / kGmStages;
            mbar_wait(FULL(s), j & 1);
            tc_fence_after();
            const uint32_t st = base + (uint32_t)s * kGmStageBytes;
            const uint64_t a_hi = make_desc(st), a_lo = make_desc(st + kGmTile);
            const uint64_t b_hi = make_desc(st + 2 * kGmTile), b_lo = make_desc(st + 3 * kGmTile);
            if (elect_one()) {
#pragma unroll
                for (int term = 0; term < 3; term++) {          // small terms first
                    const uint64_t ad = term == 0 ? a_lo : a_hi;
                    const uint64_t bd = term == 1 ? b_lo : b_hi;
#pragma unroll
                    for (int kk = 0; kk < 4; kk++)
                        tc_mma_tf32(tmem, ad + (uint64_t)((kk * 32) >> 4), bd + (uint64_t)((kk * 32) >> 4), idesc,
                                    (kc | term | kk) ? 1u : 0u);
                }
                tc_commit(EMPTY(s));
                if (kc == n_chunks - 1) tc_commit(TFULL);
            }
            __syncwarp();
        }
    } else {
        // ===== producers: row r of the A tile and row r of the Bt tile ========================================
        const int r = (warp - 5) * 32 + lane;
        const bool a_ok = m0 + r < M;
        const float4 *arow = reinterpret_cast<const float4 *>(A + (size_t)(a_ok ? m0 + r : 0) * lda);
        const float4 *brow = reinterpret_cast<const float4 *>(Bt + (size_t)(n0 + r) * ldb);
        float4 a[8], b[8];
#pragma unroll
        for (int c = 0; c < 8; c++) {
            a[c] = a_ok ? __ldg(arow + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            b[c] = __ldg(brow + c);
        }
        for (int kc = 0; kc < n_chunks; kc++) {
            const int s = kc % kGmStages, j = kc / kGmStages;
            float4 an[8], bn[8];
            if (kc + 1 < n_chunks) {                            // next chunk in flight while this one is split and stored
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    an[c] = a_ok ? __ldg(arow + (kc + 1) * 8 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
                    bn[c] = __ldg(brow + (kc + 1) * 8 + c);
                }
            }
            mbar_wait(EMPTY(s), (j & 1) ^ 1);
            const uint32_t st = base + (uint32_t)s * kGmStageBytes;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const uint32_t off = gm_off(r, c);
                const float va[4] = {a[c].x, a[c].y, a[c].z, a[c].w}, vb[4] = {b[c].x, b[c].y, b[c].z, b[c].w};
                sts_split4(st + off, st + kGmTile + off, va);
                sts_split4(st + 2 * kGmTile + off, st + 3 * kGmTile + off, vb);
            }
            fence_async_smem();                                 // generic-proxy stores -> visible to the tensor core
            mbar_arrive(FULL(s));
            if (kc + 1 < n_chunks) {
#pragma unroll
                for (int c = 0; c < 8; c++) { a[c] = an[c]; b[c] = bn[c]; }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
    }
}

// ---- the same GEMM with PRE-SPLIT, PRE-SWIZZLED weights ------------------------------------------------------------
// The weight operand does not change between calls: mgar_gemm_pack_bt splits it once into tf32 hi / lo and lays every
// (128 rows x 32 floats) tile out exactly as the MMA wants it in shared memory (128-byte swizzle), hi and lo tile back to
// back: [n_tile][k_chunk][hi 16 KB | lo 16 KB].  The kernel then brings a chunk's 32 KB of B with ONE cp.async.bulk (the TMA
// engine's bulk copy) that completes on the stage's full barrier - no registers, no split, no shared-memory stores for B -
// and the four producer warps only stream A, two K chunks ahead in registers (a K chunk's MMAs last ~0.4 us, a trip to
// L2 / HBM ~1 us: one chunk of prefetch, as the plain kernel has, leaves the tensor pipe waiting).
__global__ void __launch_bounds__(256)
gemm_pack_bt_kernel(int N, int K, const float *__restrict__ Bt, int ldb, float *__restrict__ packed) {
    // one thread per 16-byte chunk of the source: row n, chunk c16 of K chunk kc
    const long long total = (long long)N * (K / 4);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / (K / 4)), c = (int)(i % (K / 4));
        const int kc = c >> 3, c16 = c & 7, r = n & 127, nt = n >> 7;
        const float4 v = __ldg(reinterpret_cast<const float4 *>(Bt + (size_t)n * ldb) + c);
        float4 hi, lo;
        hi.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); lo.x = __fsub_rn(v.x, hi.x);
        hi.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); lo.y = __fsub_rn(v.y, hi.y);
        hi.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); lo.z = __fsub_rn(v.z, hi.z);
        hi.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); lo.w = __fsub_rn(v.w, hi.w);
        char *tile = reinterpret_cast<char *>(packed) + ((size_t)nt * (K / kGmKC) + kc) * 2 * kGmTile;
        *reinterpret_cast<float4 *>(tile + gm_off(r, c16)) = hi;
        *reinterpret_cast<float4 *>(tile + kGmTile + gm_off(r, c16)) = lo;
    }
}

constexpr uint32_t kGpStageBytes = 4 * kGmTile;            // A hi, A lo (written by the producers) | B hi, B lo (bulk copy)
constexpr int kGpStages = 3;
constexpr uint32_t kGpSmem = kGpStages * kGpStageBytes + 1024;

// Up to kGmGroups independent products of ONE shape in a launch (grid.z = group): the tiles of all of them fill the SMs
// together - a (6400 x 512) x 512 product alone is 200 tiles on 148 SMs, i.e. two waves for 1.35 waves of work.
constexpr int kGmGroups = 4;
struct GemmGroups {
    const float *A[kGmGroups], *packed[kGmGroups], *bias[kGmGroups];
    float *C[kGmGroups];
    int lda[kGmGroups], ldc[kGmGroups];
};

__global__ void __launch_bounds__(kGmThreads, 1)
gemm_tf32x3_packed_kernel(int M, int N, int K, const __grid_constant__ GemmGroups G, int relu) {
    const float *__restrict__ A = G.A[blockIdx.z];
    const float *__restrict__ packed = G.packed[blockIdx.z];
    const float *__restrict__ bias = G.bias[blockIdx.z];
    float *__restrict__ C = G.C[blockIdx.z];
    const int lda = G.lda[blockIdx.z], ldc = G.ldc[blockIdx.z];
    extern __shared__ uint8_t gm_smem[];
    __shared__ __align__(8) uint64_t s_bar[2 * kGpStages + 1];
    __shared__ uint32_t s_tmem;
    const uint32_t base = (smem_u32(gm_smem) + 1023u) & ~1023u;
    const uint32_t bar0 = smem_u32(s_bar);
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (kGpStages + s); };
    const uint32_t TFULL = bar0 + 8u * (2 * kGpStages);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.x * 128, m0 = blockIdx.y * 128;
    const int n_chunks = K / kGmKC;

    if (tid == 0) {
        for (int s = 0; s < kGpStages; s++) { mbar_init(FULL(s), 129); mbar_init(EMPTY(s), 1); }   // 128 producers + the bulk copy
        mbar_init(TFULL, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (warp < 4) {
        mbar_wait(TFULL, 0);
        tc_fence_after();
        const int m = m0 + warp * 32 + lane;
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
        for (int q = 0; q < 4; q++) {
            uint32_t v[32];
            tc_ld32(taddr + (uint32_t)(q * 32), v);
            tc_wait_ld();
            if (m < M) {
                float *out = C + (size_t)m * ldc + n0 + q * 32;
                const float *bq = bias ? bias + n0 + q * 32 : nullptr;
#pragma unroll
                for (int c = 0; c < 32; c += 4) {
                    float4 o = make_float4(__uint_as_float(v[c]), __uint_as_float(v[c + 1]), __uint_as_float(v[c + 2]),
                                           __uint_as_float(v[c + 3]));
                    if (bq) {
                        const float4 bb = __ldg(reinterpret_cast<const float4 *>(bq + c));
                        o.x = __fadd_rn(o.x, bb.x); o.y = __fadd_rn(o.y, bb.y); o.z = __fadd_rn(o.z, bb.z); o.w = __fadd_rn(o.w, bb.w);
                    }
                    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
                    *reinterpret_cast<float4 *>(out + c) = o;
                }
            }
        }
        tc_fence_before();
    } else if (warp == 4) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        for (int kc = 0; kc < n_chunks; kc++) {
            const int s = kc % kGpStages, j = kc / kGpStages;
            mbar_wait(FULL(s), j & 1);                          // A stored by the producers AND the B tiles landed
            tc_fence_after();
            const uint32_t st = base + (uint32_t)s * kGpStageBytes;
            const uint64_t a_hi = make_desc(st), a_lo = make_desc(st + kGmTile);
            const uint64_t b_hi = make_desc(st + 2 * kGmTile), b_lo = make_desc(st + 3 * kGmTile);
            if (elect_one()) {
#pragma unroll
                for (int term = 0; term < 3; term++) {
                    const uint64_t ad = term == 0 ? a_lo : a_hi;
                    const uint64_t bd = term == 1 ? b_lo : b_hi;
#pragma unroll
                    for (int kk = 0; kk < 4; kk++)
                        tc_mma_tf32(tmem, ad + (uint64_t)((kk * 32) >> 4), bd + (uint64_t)((kk * 32) >> 4), idesc,
                                    (kc | term | kk) ? 1u : 0u);
                }
                tc_commit(EMPTY(s));
                if (kc == n_chunks - 1) tc_commit(TFULL);
            }
            __syncwarp();
        }
    } else {
        const int r = (warp - 5) * 32 + lane;
        const bool a_ok = m0 + r < M;
        const float4 *arow = reinterpret_cast<const float4 *>(A + (size_t)(a_ok ? m0 + r : 0) * lda);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const char *btiles = reinterpret_cast<const char *>(packed) + (size_t)blockIdx.x * n_chunks * 2 * kGmTile;
        float4 a[3][8];                                         // chunks kc, kc + 1, kc + 2 (rotating, fully unrolled by 3)
#pragma unroll
        for (int d = 0; d < 2; d++)
#pragma unroll
            for (int c = 0; c < 8; c++) a[d][c] = (a_ok && d < n_chunks) ? __ldg(arow + d * 8 + c) : z4;
        auto step = [&](int kc, float4 (&cur)[8], float4 (&nxt2)[8]) {
            const int s = kc % kGpStages, j = kc / kGpStages;
            if (kc + 2 < n_chunks) {
#pragma unroll
                for (int c = 0; c < 8; c++) nxt2[c] = a_ok ? __ldg(arow + (kc + 2) * 8 + c) : z4;
            }
            mbar_wait(EMPTY(s), (j & 1) ^ 1);
            const uint32_t st = base + (uint32_t)s * kGpStageBytes;
            if (r == 0) {                                       // this chunk's weight tiles: one bulk copy onto the full barrier
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(FULL(s)), "r"(2 * kGmTile) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(st + 2 * kGmTile), "l"(btiles + (size_t)kc * 2 * kGmTile), "r"(2 * kGmTile), "r"(FULL(s)) : "memory");
            }
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const uint32_t off = gm_off(r, c);
                const float va[4] = {cur[c].x, cur[c].y, cur[c].z, cur[c].w};
                sts_split4(st + off, st + kGmTile + off, va);
            }
            fence_async_smem();
            mbar_arrive(FULL(s));
        };
        for (int kc = 0; kc < n_chunks; kc += 3) {
            step(kc, a[0], a[2]);
            if (kc + 1 < n_chunks) step(kc + 1, a[1], a[0]);
            if (kc + 2 < n_chunks) step(kc + 2, a[2], a[1]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
    }
}

}  // namespace
}  // namespace mgar

using namespace mgar;

extern "C" int mgar_gemm_tf32x3_supported(int M, int N, int K) {
    return M >= 1 && N >= 128 && N % 128 == 0 && K >= 32 && K % 32 == 0 && !getenv("MGAR_NO_TC_GEMM");
}

extern "C" int mgar_gemm_tf32x3(int M, int N, int K, const float *A, int lda, const float *Bt, int ldb, const float *bias,
                                int relu, float *C, int ldc, void *stream) {
    MGAR_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm_tf32x3: negative size");
    if (M == 0 || N == 0) return MGAR_OK;
    MGAR_REQUIRE(mgar_gemm_tf32x3_supported(M, N, K), "gemm_tf32x3: unsupported shape M=%d N=%d K=%d (N %% 128 == 0, K %% 32 == 0; "
                 "ask mgar_gemm_tf32x3_supported first)", M, N, K);
    MGAR_REQUIRE(A && Bt && C, "gemm_tf32x3: null pointer");
    MGAR_REQUIRE(lda >= K && ldb >= K && ldc >= N && lda % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0,
                 "gemm_tf32x3: leading dimensions must cover the rows and be multiples of 4 floats");
    MGAR_REQUIRE(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(Bt) | reinterpret_cast<uintptr_t>(C) |
                   reinterpret_cast<uintptr_t>(bias)) & 15) == 0, "gemm_tf32x3: operands must be 16-byte aligned");
    MGAR_REQUIRE((M + 127) / 128 <= 65535, "gemm_tf32x3: M=%d exceeds grid.y", M);
    MGAR_CUDA(cudaFuncSetAttribute(gemm_tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGmSmem));
    dim3 grid(N / 128, (M + 127) / 128);
    gemm_tf32x3_kernel<<<grid, kGmThreads, kGmSmem, (cudaStream_t)stream>>>(M, N, K, A, lda, Bt, ldb, bias, relu, C, ldc);
    count_launch();
    return check_launch("gemm_tf32x3");
}

extern "C" size_t mgar_gemm_packed_bytes(int N, int K) {
    return (N > 0 && K > 0 && N % 128 == 0 && K % 32 == 0) ? (size_t)N * K * 2 * sizeof(float) : 0;
}

extern "C" int mgar_gemm_pack_bt(int N, int K, const float *Bt, int ldb, float *packed, void *stream) {
    MGAR_REQUIRE(N >= 128 && N % 128 == 0 && K >= 32 && K % 32 == 0, "gemm_pack_bt: N %% 128 == 0 and K %% 32 == 0 required (N=%d K=%d)", N, K);
    MGAR_REQUIRE(Bt && packed && ldb >= K && ldb % 4 == 0, "gemm_pack_bt: bad operand");
    MGAR_REQUIRE(((reinterpret_cast<uintptr_t>(Bt) | reinterpret_cast<uintptr_t>(packed)) & 15) == 0, "gemm_pack_bt: operands must be 16-byte aligned");
    const long long total = (long long)N * (K / 4);
    int grid = div_up(total, 256);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    gemm_pack_bt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(N, K, Bt, ldb, packed);
    count_launch();
    return check_launch("gemm_pack_bt");
}

extern "C" int mgar_gemm_tf32x3_packed(int M, int N, int K, const float *A, int lda, const float *packed, const float *bias,
                                       int relu, float *C, int ldc, void *stream) {
    MGAR_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm_tf32x3_packed: negative size");
    if (M == 0 || N == 0) return MGAR_OK;
    MGAR_REQUIRE(mgar_gemm_tf32x3_supported(M, N, K), "gemm_tf32x3_packed: unsupported shape M=%d N=%d K=%d", M, N, K);
    MGAR_REQUIRE(A && packed && C, "gemm_tf32x3_packed: null pointer");
    MGAR_REQUIRE(lda >= K && ldc >= N && lda % 4 == 0 && ldc % 4 == 0, "gemm_tf32x3_packed: bad leading dimension");
    MGAR_REQUIRE(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(packed) | reinterpret_cast<uintptr_t>(C) |
                   reinterpret_cast<uintptr_t>(bias)) & 15) == 0, "gemm_tf32x3_packed: operands must be 16-byte aligned");
    MGAR_REQUIRE((M + 127) / 128 <= 65535, "gemm_tf32x3_packed: M=%d exceeds grid.y", M);
    MGAR_CUDA(cudaFuncSetAttribute(gemm_tf32x3_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGpSmem));
    dim3 grid(N / 128, (M + 127) / 128);
    GemmGroups G = {};
    G.A[0] = A; G.packed[0] = packed; G.bias[0] = bias; G.C[0] = C; G.lda[0] = lda; G.ldc[0] = ldc;
    gemm_tf32x3_packed_kernel<<<grid, kGmThreads, kGpSmem, (cudaStream_t)stream>>>(M, N, K, G, relu);
    count_launch();
    return check_launch("gemm_tf32x3_packed");
}

extern "C" int mgar_gemm_tf32x3_packed_grouped(int groups, int M, int N, int K, const float *const *A, const int *lda,
                                               const float *const *packed, const float *const *bias, int relu,
                                               float *const *C, const int *ldc, void *stream) {
    MGAR_REQUIRE(groups >= 1 && groups <= kGmGroups, "gemm_tf32x3_packed_grouped: 1 .. %d groups, got %d", kGmGroups, groups);
    MGAR_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm_tf32x3_packed_grouped: negative size");
    if (M == 0 || N == 0) return MGAR_OK;
    MGAR_REQUIRE(mgar_gemm_tf32x3_supported(M, N, K), "gemm_tf32x3_packed_grouped: unsupported shape M=%d N=%d K=%d", M, N, K);
    MGAR_REQUIRE(A && lda && packed && C && ldc, "gemm_tf32x3_packed_grouped: null pointer");
    GemmGroups G = {};
    for (int g = 0; g < groups; g++) {
        MGAR_REQUIRE(A[g] && packed[g] && C[g], "gemm_tf32x3_packed_grouped: null operand in group %d", g);
        MGAR_REQUIRE(lda[g] >= K && ldc[g] >= N && lda[g] % 4 == 0 && ldc[g] % 4 == 0, "gemm_tf32x3_packed_grouped: bad leading dimension");
        const float *bg = bias ? bias[g] : nullptr;
        MGAR_REQUIRE(((reinterpret_cast<uintptr_t>(A[g]) | reinterpret_cast<uintptr_t>(packed[g]) | reinterpret_cast<uintptr_t>(C[g]) |
                       reinterpret_cast<uintptr_t>(bg)) & 15) == 0, "gemm_tf32x3_packed_grouped: operands must be 16-byte aligned");
        G.A[g] = A[g]; G.packed[g] = packed[g]; G.bias[g] = bg; G.C[g] = C[g]; G.lda[g] = lda[g]; G.ldc[g] = ldc[g];
    }
    MGAR_REQUIRE((M + 127) / 128 <= 65535, "gemm_tf32x3_packed_grouped: M=%d exceeds grid.y", M);
    MGAR_CUDA(cudaFuncSetAttribute(gemm_tf32x3_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGpSmem));
    dim3 grid(N / 128, (M + 127) / 128, groups);
    gemm_tf32x3_packed_kernel<<<grid, kGmThreads, kGpSmem, (cudaStream_t)stream>>>(M, N, K, G, relu);
    count_launch();
    return check_launch("gemm_tf32x3_packed_grouped");
}
