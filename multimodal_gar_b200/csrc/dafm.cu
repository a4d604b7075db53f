// Distance Aware Fusion Module, attention core: one CTA per (frame, stream).
//
// Reference: FusionAttention_mat.forward, model/gat_model.py:466-511, applied to one frame at a time inside a Python
// loop (:1396-1475).  Per stream it runs, as separate library kernels on (n,512) / (n,n) tensors with n <= 100 persons:
//   Att = softmax((Q K^T) * E / sqrt(D), dim=1) ; X = LayerNorm(Att V + residual)          (:485-489 / :501-505)
// followed by X = LayerNorm(X + FFN(X)) (:490-491 / :506-507).  The (B*n,512)x(512,512) projections and FFN linears are
// plain library GEMMs; what is fused here is everything between them:
//   dafm_attention_kernel : CTA per (frame, stream, block of 64 query rows), 2-3 CTAs per SM; scores (fp32 FFMA, 4x8
//                           register tiles, operands staged in shared memory in 32-wide D chunks and read as LDS.128),
//                           * E * scale, key mask, row softmax, Att V + residual, LayerNorm - the n x n map never
//                           leaves shared memory and the pre-LayerNorm rows are re-read by the CTA that wrote them;
//   dafm_add_ln_kernel    : LayerNorm(x + y) with padded persons zeroed (warp per row, the row in registers).
// Both streams of a module (RGB <- LiDAR queries, LiDAR <- RGB queries) go in ONE launch (grid.y = stream).
#include "mgar_common.cuh"

namespace mgar {

constexpr int kDaThreads = 256;
constexpr int kDaChunk = 32;         // D columns per shared-memory stage
constexpr int kDaMaxN = 128;         // persons per frame (the reference pads to 100)
constexpr int kDaMaxStreams = 2;
constexpr int kDaLd = kDaChunk + 4;  // operand row stride: 16-byte aligned rows, conflict-free 128-bit loads per quarter warp
__host__ __device__ constexpr int da_row(int n) { return ((n + 3) & ~3) + 4; }

struct DafmStreams {
    const float *q[kDaMaxStreams], *k[kDaMaxStreams], *v[kDaMaxStreams], *resid[kDaMaxStreams];
    const float *gamma[kDaMaxStreams], *beta[kDaMaxStreams];
    float *out[kDaMaxStreams];
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// LayerNorm of one row held as kPer values per lane (element c = lane + 32 * t): biased variance, eps inside the
// square root, as torch.nn.LayerNorm.
template <int kPer>
__device__ __forceinline__ void ln_row(float (&x)[kPer], int D, float eps, const float *__restrict__ gamma,
                                       const float *__restrict__ beta, float *__restrict__ dst, int lane) {
    float s = 0.f;
#pragma unroll
    for (int t = 0; t < kPer; ++t) s += (lane + 32 * t < D) ? x[t] : 0.f;
    const float mean = warp_sum(s) / (float)D;
    float q = 0.f;
#pragma unroll
    for (int t = 0; t < kPer; ++t) {
        const float d = x[t] - mean;
        q += (lane + 32 * t < D) ? d * d : 0.f;
    }
    const float rstd = 1.0f / sqrtf(warp_sum(q) / (float)D + eps);
#pragma unroll
    for (int t = 0; t < kPer; ++t) {
        const int c = lane + 32 * t;
        if (c < D) dst[c] = (x[t] - mean) * rstd * __ldg(gamma + c) + __ldg(beta + c);
    }
}

constexpr int kLnPer = 16;           // D <= 512

constexpr int kDaRows = 64;          // query rows per CTA: grid.z = ceil(n / 64) row blocks
constexpr int kDaVChunk = 64;        // D columns per V stage (two per lane)
constexpr size_t da_smem_floats(int n) {
    return (size_t)kDaRows * da_row(n) + (size_t)kDaRows * kDaLd + (size_t)kDaMaxN * kDaVChunk;
}

__global__ void __launch_bounds__(kDaThreads, 2)
dafm_attention_kernel(int n, int D, int ld, DafmStreams P, const float *__restrict__ E, const uint8_t *__restrict__ valid,
                      float scale, float eps) {
    extern __shared__ float smem_da[];
    const int b = blockIdx.x, s = blockIdx.y, tid = threadIdx.x;
    const int r0 = blockIdx.z * kDaRows, rows = min(kDaRows, n - r0);   // this CTA's query rows
    const int np = da_row(n);                                    // row stride of the rows x n map (multiple of 4 floats)
    float *S = smem_da;                                          // [64][np], columns n..n4 are zero after the softmax
    float *Qs = S + (size_t)kDaRows * np;                        // [64][36]   (rows >= rows are zero)
    float *Ks = Qs + kDaRows * kDaLd;                            // [128][36]  (rows >= n are zero); later the V stage [n4][64]
    const float *q = P.q[s] + ((size_t)b * n + r0) * ld, *k = P.k[s] + (size_t)b * n * ld, *v = P.v[s] + (size_t)b * n * ld;
    const float *resid = P.resid[s] + ((size_t)b * n + r0) * D;
    float *out = P.out[s] + ((size_t)b * n + r0) * D;

    // ---- scores: thread (ty, tx) owns rows ty + 16 a (a < 4), columns tx + 16 c (c < 8)
    const int tx = tid & 15, ty = tid >> 4;
    float acc[4][8];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[a][c] = 0.f;
    // operand staging: 128-bit loads, all of a chunk's loads in flight together, and the NEXT chunk's loads issued
    // before the current chunk's FFMAs (registers kq / kk4) so their latency hides behind the arithmetic
    float4 kq[4], qq[2];
    auto load_qk = [&](int d0) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int e = tid + u * kDaThreads, r = e >> 3, c = (e & 7) * 4;
            kq[u] = (r < n && d0 + c < D) ? __ldg(reinterpret_cast<const float4 *>(k + (size_t)r * ld + d0 + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int e = tid + u * kDaThreads, r = e >> 3, c = (e & 7) * 4;
            qq[u] = (r < rows && d0 + c < D) ? __ldg(reinterpret_cast<const float4 *>(q + (size_t)r * ld + d0 + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    load_qk(0);
    for (int d0 = 0; d0 < D; d0 += kDaChunk) {
        __syncthreads();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int e = tid + u * kDaThreads;
            *reinterpret_cast<float4 *>(Ks + (e >> 3) * kDaLd + (e & 7) * 4) = kq[u];
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int e = tid + u * kDaThreads;
            *reinterpret_cast<float4 *>(Qs + (e >> 3) * kDaLd + (e & 7) * 4) = qq[u];
        }
        __syncthreads();
        if (d0 + kDaChunk < D) load_qk(d0 + kDaChunk);
#pragma unroll 2
        for (int kk = 0; kk < kDaChunk; kk += 4) {
            float4 qa[4], kc[8];
#pragma unroll
            for (int a = 0; a < 4; ++a) qa[a] = *reinterpret_cast<const float4 *>(Qs + (ty + 16 * a) * kDaLd + kk);
#pragma unroll
            for (int c = 0; c < 8; ++c) kc[c] = *reinterpret_cast<const float4 *>(Ks + (tx + 16 * c) * kDaLd + kk);
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    acc[a][c] = fmaf(qa[a].x, kc[c].x, acc[a][c]);
                    acc[a][c] = fmaf(qa[a].y, kc[c].y, acc[a][c]);
                    acc[a][c] = fmaf(qa[a].z, kc[c].z, acc[a][c]);
                    acc[a][c] = fmaf(qa[a].w, kc[c].w, acc[a][c]);
                }
        }
    }
    const float *Eb = E + ((size_t)b * n + r0) * n;
    float ev[4][8];                                              // all 32 loads in flight before the first use
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int i = ty + 16 * a, j = tx + 16 * c;
            ev[a][c] = (i < rows && j < n) ? __ldg(Eb + (size_t)i * n + j) : 0.f;
        }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int i = ty + 16 * a, j = tx + 16 * c;
            if (i < rows && j < n) S[i * np + j] = acc[a][c] * ev[a][c] / scale;   // (QK^T * E) / sqrt(D), :485
        }
    __syncthreads();

    // ---- row softmax over the valid keys (warp per row)
    const int lane = tid & 31, warp = tid >> 5;
    const uint8_t *vb = valid ? valid + (size_t)b * n : nullptr;
    for (int i = warp; i < rows; i += kDaThreads / 32) {
        float x[kDaMaxN / 32];
        float m = -INFINITY;
#pragma unroll
        for (int t = 0; t < kDaMaxN / 32; ++t) {
            const int j = lane + 32 * t;
            x[t] = (j < n && (!vb || vb[j])) ? S[i * np + j] : -INFINITY;
            m = fmaxf(m, x[t]);
        }
        m = warp_max(m);
        float sum = 0.f;
#pragma unroll
        for (int t = 0; t < kDaMaxN / 32; ++t) {
            x[t] = expf(x[t] - m);
            sum += x[t];
        }
        sum = warp_sum(sum);
#pragma unroll
        for (int t = 0; t < kDaMaxN / 32; ++t) {
            const int j = lane + 32 * t;
            if (j < np) S[i * np + j] = j < n ? x[t] / sum : 0.f;      // pad columns are zero: the AV loop reads whole float4s
        }
    }

    // ---- Att V + residual, 64 columns at a time: warp w owns rows w + 8 t (t < 8), lane l columns l and l + 32
    float *Vs = Ks;
    const int n4 = (n + 3) & ~3;
    float4 vq[8];
    auto load_v = [&](int d0) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int e = tid + u * kDaThreads, r = e >> 4, c = (e & 15) * 4;
            vq[u] = (r < n && d0 + c < D) ? __ldg(reinterpret_cast<const float4 *>(v + (size_t)r * ld + d0 + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    load_v(0);
    for (int d0 = 0; d0 < D; d0 += kDaVChunk) {
        __syncthreads();
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int e = tid + u * kDaThreads;
            if ((e >> 4) < n4) *reinterpret_cast<float4 *>(Vs + (e >> 4) * kDaVChunk + (e & 15) * 4) = vq[u];
        }
        __syncthreads();
        if (d0 + kDaVChunk < D) load_v(d0 + kDaVChunk);
        float o[kDaRows / 8][2];                                 // starts from the residual: its loads overlap the FFMAs
#pragma unroll
        for (int t = 0; t < kDaRows / 8; ++t)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int i = warp + 8 * t, c = d0 + 32 * h + lane;
                o[t][h] = (i < rows && c < D) ? __ldg(resid + (size_t)i * D + c) : 0.f;
            }
        for (int j = 0; j < n4; j += 4) {
            float va[4], vb2[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                va[u] = Vs[(j + u) * kDaVChunk + lane];
                vb2[u] = Vs[(j + u) * kDaVChunk + 32 + lane];
            }
#pragma unroll
            for (int t = 0; t < kDaRows / 8; ++t) {
                const float4 a = *reinterpret_cast<const float4 *>(S + (warp + 8 * t) * np + j);   // one broadcast LDS.128 per warp
                o[t][0] = fmaf(a.x, va[0], o[t][0]); o[t][1] = fmaf(a.x, vb2[0], o[t][1]);
                o[t][0] = fmaf(a.y, va[1], o[t][0]); o[t][1] = fmaf(a.y, vb2[1], o[t][1]);
                o[t][0] = fmaf(a.z, va[2], o[t][0]); o[t][1] = fmaf(a.z, vb2[2], o[t][1]);
                o[t][0] = fmaf(a.w, va[3], o[t][0]); o[t][1] = fmaf(a.w, vb2[3], o[t][1]);
            }
        }
#pragma unroll
        for (int t = 0; t < kDaRows / 8; ++t) {
            const int i = warp + 8 * t;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = d0 + 32 * h + lane;
                if (i < rows && c < D) out[(size_t)i * D + c] = o[t][h];
            }
        }
    }
    __syncthreads();                                             // the CTA's own global writes are visible to it

    // ---- LayerNorm in place (warp per row)
    for (int i = warp; i < rows; i += kDaThreads / 32) {
        float x[kLnPer];
#pragma unroll
        for (int t = 0; t < kLnPer; ++t) x[t] = (lane + 32 * t < D) ? out[(size_t)i * D + lane + 32 * t] : 0.f;
        ln_row<kLnPer>(x, D, eps, P.gamma[s], P.beta[s], out + (size_t)i * D, lane);
    }
}

// out = LayerNorm(x + y) row-wise; rows whose valid byte is 0 are written as zeros.
__global__ void __launch_bounds__(256)
dafm_add_ln_kernel(int rows, int D, const float *__restrict__ x, const float *__restrict__ y, const float *__restrict__ gamma,
                   const float *__restrict__ beta, const uint8_t *__restrict__ valid, float eps, float *__restrict__ out) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    float v[kLnPer];
    if (valid && !valid[row]) {
        for (int c = lane; c < D; c += 32) out[(size_t)row * D + c] = 0.f;
        return;
    }
#pragma unroll
    for (int t = 0; t < kLnPer; ++t) {
        const int c = lane + 32 * t;
        v[t] = c < D ? __ldg(x + (size_t)row * D + c) + __ldg(y + (size_t)row * D + c) : 0.f;
    }
    ln_row<kLnPer>(v, D, eps, gamma, beta, out + (size_t)row * D, lane);
}

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_dafm_attention(int B, int n, int D, int n_streams, int ld, const float *const *q, const float *const *k,
                                   const float *const *v, const float *const *resid, const float *E,
                                   const unsigned char *valid, const float *const *gamma, const float *const *beta,
                                   float eps, float *const *out, void *stream) {
    MGAR_REQUIRE(B >= 0 && n >= 1 && n <= kDaMaxN, "dafm_attention: n=%d persons per frame outside [1, %d]", n, kDaMaxN);
    MGAR_REQUIRE(D >= 1 && D <= 32 * kLnPer && ld >= D, "dafm_attention: D=%d (ld=%d) outside [1, %d]", D, ld, 32 * kLnPer);
    MGAR_REQUIRE(n_streams >= 1 && n_streams <= kDaMaxStreams, "dafm_attention: n_streams must be 1 or 2");
    MGAR_REQUIRE(D % 4 == 0 && ld % 4 == 0, "dafm_attention: D=%d and ld=%d must be multiples of 4 (128-bit operand loads)", D, ld);
    MGAR_REQUIRE(B <= 0x7fffffff / (n * n), "dafm_attention: batch too large");
    if (B == 0) return MGAR_OK;
    MGAR_REQUIRE(q && k && v && resid && E && gamma && beta && out, "dafm_attention: null pointer");
    DafmStreams P;
    for (int s = 0; s < kDaMaxStreams; ++s) {
        const int t = s < n_streams ? s : 0;
        MGAR_REQUIRE(q[t] && k[t] && v[t] && resid[t] && gamma[t] && beta[t] && out[t], "dafm_attention: null pointer (stream %d)", t);
        MGAR_REQUIRE(((reinterpret_cast<uintptr_t>(q[t]) | reinterpret_cast<uintptr_t>(k[t]) | reinterpret_cast<uintptr_t>(v[t])) & 15) == 0,
                     "dafm_attention: q / k / v must be 16-byte aligned (stream %d)", t);
        P.q[s] = q[t]; P.k[s] = k[t]; P.v[s] = v[t]; P.resid[s] = resid[t];
        P.gamma[s] = gamma[t]; P.beta[s] = beta[t]; P.out[s] = out[t];
    }
    const size_t smem = da_smem_floats(n) * sizeof(float);
    // per call: the attribute belongs to the current device's context (a process may drive several GPUs)
    MGAR_CUDA(cudaFuncSetAttribute(dafm_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(da_smem_floats(kDaMaxN) * sizeof(float))));
    dafm_attention_kernel<<<dim3(B, n_streams, div_up(n, kDaRows)), kDaThreads, smem, (cudaStream_t)stream>>>(n, D, ld, P, E, valid,
                                                                                           sqrtf((float)D), eps);
    count_launch();
    return check_launch("dafm_attention");
}

extern "C" int mgar_dafm_add_layernorm(int rows, int D, const float *x, const float *y, const float *gamma, const float *beta,
                                       const unsigned char *valid, float eps, float *out, void *stream) {
    MGAR_REQUIRE(rows >= 0 && D >= 1 && D <= 32 * kLnPer, "dafm_add_layernorm: D=%d outside [1, %d]", D, 32 * kLnPer);
    if (rows == 0) return MGAR_OK;
    MGAR_REQUIRE(x && y && gamma && beta && out, "dafm_add_layernorm: null pointer");
    dafm_add_ln_kernel<<<div_up(rows, 8), 256, 0, (cudaStream_t)stream>>>(rows, D, x, y, gamma, beta, valid, eps, out);
    count_launch();
    return check_launch("dafm_add_layernorm");
}
