// Set-abstraction shared MLP in TRAINING mode (BatchNorm with batch statistics) and its backward pass.
//
// Reference: StackSAModuleMSG.forward in train() (pointnet2_stack/pointnet2_modules.py:78-112): QueryAndGroup
// (pointnet2_utils.py:123-159) -> [Conv2d 1x1 (no bias), BatchNorm2d (batch statistics over ALL M*nsample
// positions, biased variance, eps), ReLU] x L -> max/avg pool over the samples; backward through cuDNN / ATen
// autograd and GroupingOperation.backward (group_points_gpu.cu:15-44: atomicAdd scatter to the features).
//
// Batch statistics couple every position, so the layers cannot be fused into one pass.  The pre-activations
// z_l (P x C_l, P = M*nsample) are MATERIALISED in a caller-provided workspace (sized by
// mgar_sa_mlp_train_workspace_bytes; the 180 GB of HBM3e hold a whole training batch) and kept for the
// backward pass:
//   forward :  x0 = gather | for l: z_l = act_{l-1}(z_{l-1}) . W_l^T, column sums (fp64 atomics) -> mean, invstd |
//              out = pool_k relu(bn_L(z_L)) (+ arg-max for the backward)
//   backward:  for l = L..1: g_y = g_h * (y > 0); s1 = sum g_y, s2 = sum g_y*xhat (fp64) -> dgamma, dbeta;
//              g_z = gamma*invstd*(g_y - s1/P - xhat*s2/P); dW_l = a_{l-1}^T g_z (register accumulators, one
//              atomic flush per CTA); g_h_{l-1} = g_z W_l; at l = 1 the feature columns are scattered to
//              grad_features with atomicAdd.  Gradients w.r.t. xyz / new_xyz are not produced (the branch's
//              coordinates never require grad).
// Exact-fp32 FFMA arithmetic; this path favours clarity over speed (it is not the inference hot path).
#include "mgar_common.cuh"
#include "stack_prefix.cuh"
#include "sa_mlp_train_fast.cuh"

namespace mgar {

namespace {

constexpr int kTrMaxLayers = 6;
constexpr int kTrThreads = 256;

struct TrainSpec {
    int L, K, M;
    long long P;                       // M * K positions
    int ch[kTrMaxLayers + 1];
    int w_off[kTrMaxLayers];           // float offset of W_l^T [ch[l]][ch[l+1]] in the weight block
    int g_off[kTrMaxLayers];           // float offset of gamma_l (then beta_l) in the gamma/beta block
    // workspace offsets in BYTES
    size_t x0, z[kTrMaxLayers], argmax, stat64, stat32, ss32, gbuf[2], bstat64, coef, total;
    int c_off[kTrMaxLayers];           // 2 * (channels of the layers before l): offset of layer l in stat64 / stat32 / ss32
    int maxc;
};

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

int make_spec(TrainSpec &s, int M, int K, int L, const int *ch) {
    if (L < 1 || L > kTrMaxLayers || M < 0 || K < 1 || !ch) return -1;
    s.L = L; s.K = K; s.M = M; s.P = (long long)M * K;
    int woff = 0, goff = 0, sumc = 0, maxc = 1;
    for (int l = 0; l <= L; l++) { if (ch[l] < 1) return -1; s.ch[l] = ch[l]; }
    for (int l = 0; l < L; l++) {
        s.w_off[l] = woff; woff += ch[l] * ch[l + 1];
        s.g_off[l] = goff; goff += 2 * ch[l + 1];
        s.c_off[l] = 2 * sumc;
        sumc += ch[l + 1];
        if (ch[l + 1] > maxc) maxc = ch[l + 1];
    }
    if (ch[0] > maxc) maxc = ch[0];
    s.maxc = maxc;
    size_t off = 0;
    s.x0 = off; off = align_up(off + (size_t)s.P * ch[0] * 4);
    for (int l = 0; l < L; l++) { s.z[l] = off; off = align_up(off + (size_t)s.P * ch[l + 1] * 4); }
    s.argmax = off; off = align_up(off + (size_t)M * ch[L] * 4);
    s.stat64 = off; off = align_up(off + (size_t)2 * sumc * 8);          // per layer: sum[C], sumsq[C] (double)
    s.stat32 = off; off = align_up(off + (size_t)2 * sumc * 4);          // per layer: mean[C], invstd[C] (float)
    s.ss32 = off; off = align_up(off + (size_t)2 * sumc * 4);            // per layer: folded scale[C], shift[C]
    // backward: g_h buffers of layers L-1 .. 0 (ping-pong); the last layer's g_h comes from grad_out on the fly
    size_t gmax = 0;
    for (int l = 0; l < L; l++) { const size_t b = (size_t)s.P * ch[l] * 4; if (b > gmax) gmax = b; }
    s.gbuf[0] = off; off = align_up(off + gmax);
    s.gbuf[1] = off; off = align_up(off + gmax);
    s.bstat64 = off; off = align_up(off + (size_t)2 * maxc * 8);
    s.coef = off; off = align_up(off + (size_t)3 * maxc * 4);
    s.total = off;
    return 0;
}

__device__ __forceinline__ int find_batch_tr(const int *__restrict__ pre, int B, int v) {
    int lo = 0, hi = B;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(pre + mid) <= v) lo = mid; else hi = mid;
    }
    return lo;
}

// x0[p, :] = (xyz[src] - centre | features[src]) or zeros for empty balls (pointnet2_utils.py:146-150); warp per centroid
__global__ void __launch_bounds__(kTrThreads)
tr_group_kernel(int B, int M, int K, int C_in, const float *__restrict__ xyz, const float *__restrict__ new_xyz,
                const float *__restrict__ features, const int *__restrict__ pre_xyz, const int *__restrict__ pre_new,
                int uniform, int n_u, int m_u, const int *__restrict__ idx, const unsigned char *__restrict__ empty_mask,
                float *__restrict__ x0) {
    const int lane = threadIdx.x & 31;
    const int C0 = 3 + C_in;
    for (int m = blockIdx.x * (kTrThreads / 32) + (threadIdx.x >> 5); m < M; m += gridDim.x * (kTrThreads / 32)) {
        const long long start = uniform ? (long long)(m / m_u) * n_u : __ldg(pre_xyz + find_batch_tr(pre_new, B, m));
        const bool empty = (__ldg(idx + (size_t)m * K) < 0) || (empty_mask && empty_mask[m]);
        const float qx = __ldg(new_xyz + (size_t)m * 3), qy = __ldg(new_xyz + (size_t)m * 3 + 1), qz = __ldg(new_xyz + (size_t)m * 3 + 2);
        for (int s = lane; s < K; s += 32) {
            float *row = x0 + ((size_t)m * K + s) * C0;
            if (!empty) {
                const long long p = start + __ldg(idx + (size_t)m * K + s);
                row[0] = __fsub_rn(__ldg(xyz + p * 3), qx);
                row[1] = __fsub_rn(__ldg(xyz + p * 3 + 1), qy);
                row[2] = __fsub_rn(__ldg(xyz + p * 3 + 2), qz);
                for (int c = 0; c < C_in; c++) row[3 + c] = __ldg(features + p * C_in + c);
            } else {
                for (int c = 0; c < C0; c++) row[c] = 0.f;
            }
        }
    }
}

// Z[p, o] = sum_c act(A[p, c]) * Wt[c][o]; act = identity (sc == nullptr) or relu(a * sc[c] + sh[c]) (the previous
// layer's BatchNorm + ReLU applied on load).  CTA = TP positions staged in shared memory.
__global__ void __launch_bounds__(kTrThreads)
tr_linear_kernel(long long P, int Ci, int Co, int TP, const float *__restrict__ A, const float *__restrict__ sc,
                 const float *__restrict__ sh, const float *__restrict__ Wt, float *__restrict__ Z) {
    extern __shared__ float tr_sm[];
    float *sA = tr_sm;                                  // [TP][Ci + 1]
    const int ld = Ci + 1;
    for (long long p0 = (long long)blockIdx.x * TP; p0 < P; p0 += (long long)gridDim.x * TP) {
        const int np = (int)min((long long)TP, P - p0);
        __syncthreads();
        for (int e = threadIdx.x; e < np * Ci; e += kTrThreads) {
            const int p = e / Ci, c = e - p * Ci;
            float a = __ldg(A + (p0 + p) * Ci + c);
            if (sc) a = fmaxf(__fmaf_rn(a, __ldg(sc + c), __ldg(sh + c)), 0.f);
            sA[p * ld + c] = a;
        }
        __syncthreads();
        for (int e = threadIdx.x; e < np * Co; e += kTrThreads) {
            const int p = e / Co, o = e - p * Co;
            const float *a = sA + p * ld;
            float acc = 0.f;
            for (int c = 0; c < Ci; c++) acc = __fmaf_rn(a[c], __ldg(Wt + (size_t)c * Co + o), acc);
            Z[(p0 + p) * Co + o] = acc;
        }
    }
}

// column sums of Z and Z^2 in fp64 (one atomic pair per thread-owned column and CTA)
__global__ void __launch_bounds__(kTrThreads)
tr_colstats_kernel(long long P, int C, int rows_per_cta, const float *__restrict__ Z, double *__restrict__ sum,
                   double *__restrict__ sq) {
    const int lanes_c = C < kTrThreads ? C : kTrThreads, rpi = kTrThreads / lanes_c;
    const int my_r = threadIdx.x / lanes_c, my_c0 = threadIdx.x % lanes_c;
    if (my_r >= rpi) return;
    const long long r0 = (long long)blockIdx.x * rows_per_cta, r1 = min(P, r0 + rows_per_cta);
    for (int c = my_c0; c < C; c += lanes_c) {
        double s = 0.0, q = 0.0;
        for (long long r = r0 + my_r; r < r1; r += rpi) {
            const double v = (double)__ldg(Z + r * C + c);
            s += v; q += v * v;
        }
        atomicAdd(sum + c, s);
        atomicAdd(sq + c, q);
    }
}

// mean, biased variance (for the caller's running statistics) and the folded scale/shift of BatchNorm
__global__ void tr_finalize_kernel(long long P, int C, float eps, const double *__restrict__ sum, const double *__restrict__ sq,
                                   const float *__restrict__ gamma, const float *__restrict__ beta,
                                   float *__restrict__ mean_invstd, float *__restrict__ scale_shift,
                                   float *__restrict__ batch_stats) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const double mean = sum[c] / (double)P;
    double var = sq[c] / (double)P - mean * mean;
    if (var < 0.0) var = 0.0;
    const double invstd = 1.0 / sqrt(var + (double)eps);
    mean_invstd[c] = (float)mean;
    mean_invstd[C + c] = (float)invstd;
    const double scl = (double)gamma[c] * invstd;
    scale_shift[c] = (float)scl;
    scale_shift[C + c] = (float)((double)beta[c] - mean * scl);
    if (batch_stats) { batch_stats[c] = (float)mean; batch_stats[C + c] = (float)var; }
}

// out[m, c] = pool_k relu(z[m, k, c] * sc[c] + sh[c]); arg[m, c] = first maximising k (max-pool)
__global__ void __launch_bounds__(kTrThreads)
tr_pool_kernel(int M, int K, int C, const float *__restrict__ Z, const float *__restrict__ sc, const float *__restrict__ sh,
               float *__restrict__ out, int ld_out, int *__restrict__ arg, int pool_avg) {
    const long long e = (long long)blockIdx.x * kTrThreads + threadIdx.x;
    if (e >= (long long)M * C) return;
    const int m = (int)(e / C), c = (int)(e - (long long)m * C);
    const float a = __ldg(sc + c), b = __ldg(sh + c);
    float best = -1.f, acc = 0.f;
    int bk = 0;
    for (int k = 0; k < K; k++) {
        const float y = fmaxf(__fmaf_rn(__ldg(Z + ((size_t)m * K + k) * C + c), a, b), 0.f);
        acc += y;
        if (y > best) { best = y; bk = k; }
    }
    out[(size_t)m * ld_out + c] = pool_avg ? acc / (float)K : best;
    arg[e] = bk;
}

// ---- backward ----------------------------------------------------------------------------------------------
// gradient w.r.t. h_l = relu(bn(z_l)) at (p, c): from grad_out through the pool (last layer) or from the buffer
struct GradSrc {
    const float *buf;        // (P, C) g_h of an inner layer, or nullptr for the last layer
    const float *grad_out;   // (M, C) with row stride ld
    const int *arg;          // (M, C) arg-max sample
    int ld, K, pool_avg;
};
__device__ __forceinline__ float grad_h(const GradSrc &g, long long p, int c, int C) {
    if (g.buf) return __ldg(g.buf + p * C + c);
    const int m = (int)(p / g.K), k = (int)(p - (long long)m * g.K);
    const float go = __ldg(g.grad_out + (size_t)m * g.ld + c);
    if (g.pool_avg) return go / (float)g.K;
    return __ldg(g.arg + (size_t)m * C + c) == k ? go : 0.f;
}

// s1[c] = sum_p g_y, s2[c] = sum_p g_y * xhat  (g_y = g_h where bn(z) > 0)
__global__ void __launch_bounds__(kTrThreads)
tr_bwd_stats_kernel(long long P, int C, int rows_per_cta, const float *__restrict__ Z, const float *__restrict__ mean_invstd,
                    const float *__restrict__ scale_shift, GradSrc gs, double *__restrict__ s1, double *__restrict__ s2) {
    const int lanes_c = C < kTrThreads ? C : kTrThreads, rpi = kTrThreads / lanes_c;
    const int my_r = threadIdx.x / lanes_c, my_c0 = threadIdx.x % lanes_c;
    if (my_r >= rpi) return;
    const long long r0 = (long long)blockIdx.x * rows_per_cta, r1 = min(P, r0 + rows_per_cta);
    for (int c = my_c0; c < C; c += lanes_c) {
        const float mean = __ldg(mean_invstd + c), inv = __ldg(mean_invstd + C + c);
        const float sc = __ldg(scale_shift + c), sh = __ldg(scale_shift + C + c);
        double a = 0.0, b = 0.0;
        for (long long r = r0 + my_r; r < r1; r += rpi) {
            const float z = __ldg(Z + r * C + c);
            if (__fmaf_rn(z, sc, sh) > 0.f) {
                const float g = grad_h(gs, r, c, C);
                a += (double)g;
                b += (double)g * (double)((z - mean) * inv);
            }
        }
        atomicAdd(s1 + c, a);
        atomicAdd(s2 + c, b);
    }
}

// coef[0][c] = gamma*invstd, coef[1][c] = s1/P, coef[2][c] = s2/P; dgamma = s2, dbeta = s1
__global__ void tr_bwd_coef_kernel(long long P, int C, const double *__restrict__ s1, const double *__restrict__ s2,
                                   const float *__restrict__ gamma, const float *__restrict__ mean_invstd,
                                   float *__restrict__ coef, float *__restrict__ dgamma, float *__restrict__ dbeta) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    coef[c] = gamma[c] * mean_invstd[C + c];
    coef[C + c] = (float)(s1[c] / (double)P);
    coef[2 * C + c] = (float)(s2[c] / (double)P);
    dgamma[c] = (float)s2[c];
    dbeta[c] = (float)s1[c];
}

constexpr int kTrTP = 32;          // positions per tile of the backward layer kernel
constexpr int kTrAcc = 32;         // dW accumulators per thread: a CTA column (blockIdx.y) owns 256*32 (c_in, o) pairs

struct ScatterArgs {               // layer 1 only: where the feature columns of g_x0 go
    float *grad_features; int C_in, B, uniform, n_u, m_u;
    const int *pre_xyz, *pre_new, *idx; const unsigned char *empty_mask;
};

// One layer of the backward pass over a persistent grid: g_z tile and a_{l-1} tile in shared memory,
// dW^T[c_in][o] += sum_p a[p][c_in] g_z[p][o] in registers, g_h_{l-1} = g_z W (CTA column 0 only).
__global__ void __launch_bounds__(kTrThreads)
tr_bwd_layer_kernel(long long P, int Ci, int Co, const float *__restrict__ Z, const float *__restrict__ mean_invstd,
                    const float *__restrict__ scale_shift, const float *__restrict__ coef, GradSrc gs,
                    const float *__restrict__ Aprev, const float *__restrict__ sc_prev, const float *__restrict__ sh_prev,
                    const float *__restrict__ Wt, int w_in_smem, float *__restrict__ dWt, float *__restrict__ g_prev,
                    ScatterArgs sa) {
    extern __shared__ float tr_sm[];
    const int ldg_ = Co + 1, lda = Ci + 1;
    float *sG = tr_sm;                               // [TP][Co + 1]
    float *sA = sG + kTrTP * ldg_;                   // [TP][Ci + 1]
    float *sW = sA + kTrTP * lda;                    // [Ci][Co + 1] (optional)
    if (w_in_smem) {
        for (int e = threadIdx.x; e < Ci * Co; e += kTrThreads) sW[(e / Co) * ldg_ + (e % Co)] = __ldg(Wt + e);
    }
    float acc[kTrAcc];
#pragma unroll
    for (int j = 0; j < kTrAcc; j++) acc[j] = 0.f;
    const long long pair0 = (long long)blockIdx.y * kTrThreads * kTrAcc;
    const long long n_pairs = (long long)Ci * Co;

    for (long long p0 = (long long)blockIdx.x * kTrTP; p0 < P; p0 += (long long)gridDim.x * kTrTP) {
        const int np = (int)min((long long)kTrTP, P - p0);
        __syncthreads();
        for (int e = threadIdx.x; e < kTrTP * Co; e += kTrThreads) {
            const int p = e / Co, o = e - p * Co;
            float gz = 0.f;
            if (p < np) {
                const float z = __ldg(Z + (p0 + p) * Co + o);
                const float gy = __fmaf_rn(z, __ldg(scale_shift + o), __ldg(scale_shift + Co + o)) > 0.f ? grad_h(gs, p0 + p, o, Co) : 0.f;
                const float xh = (z - __ldg(mean_invstd + o)) * __ldg(mean_invstd + Co + o);
                gz = __ldg(coef + o) * (gy - __ldg(coef + Co + o) - xh * __ldg(coef + 2 * Co + o));
            }
            sG[p * ldg_ + o] = gz;
        }
        for (int e = threadIdx.x; e < kTrTP * Ci; e += kTrThreads) {
            const int p = e / Ci, c = e - p * Ci;
            float a = 0.f;
            if (p < np) {
                a = __ldg(Aprev + (p0 + p) * Ci + c);
                if (sc_prev) a = fmaxf(__fmaf_rn(a, __ldg(sc_prev + c), __ldg(sh_prev + c)), 0.f);
            }
            sA[p * lda + c] = a;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < kTrAcc; j++) {
            const long long e = pair0 + threadIdx.x + (long long)j * kTrThreads;
            if (e < n_pairs) {
                const int ci = (int)(e / Co), o = (int)(e - (long long)ci * Co);
                float s = 0.f;
#pragma unroll 8
                for (int p = 0; p < kTrTP; p++) s = __fmaf_rn(sA[p * lda + ci], sG[p * ldg_ + o], s);
                acc[j] += s;
            }
        }
        if (blockIdx.y == 0) {
            for (int e = threadIdx.x; e < np * Ci; e += kTrThreads) {
                const int p = e / Ci, ci = e - p * Ci;
                const float *g = sG + p * ldg_;
                float s = 0.f;
                if (w_in_smem) { const float *w = sW + ci * ldg_; for (int o = 0; o < Co; o++) s = __fmaf_rn(g[o], w[o], s); }
                else { const float *w = Wt + (size_t)ci * Co; for (int o = 0; o < Co; o++) s = __fmaf_rn(g[o], __ldg(w + o), s); }
                if (g_prev) {
                    g_prev[(p0 + p) * Ci + ci] = s;
                } else if (sa.grad_features && ci >= 3) {      // layer 1: scatter the feature columns
                    const long long pp = p0 + p;
                    const int m = (int)(pp / gs.K);
                    const bool empty = (__ldg(sa.idx + (size_t)m * gs.K) < 0) || (sa.empty_mask && sa.empty_mask[m]);
                    if (!empty) {
                        const long long start = sa.uniform ? (long long)(m / sa.m_u) * sa.n_u
                                                           : __ldg(sa.pre_xyz + find_batch_tr(sa.pre_new, sa.B, m));
                        atomicAdd(sa.grad_features + (start + __ldg(sa.idx + pp)) * sa.C_in + (ci - 3), s);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kTrAcc; j++) {
        const long long e = pair0 + threadIdx.x + (long long)j * kTrThreads;
        if (e < n_pairs) atomicAdd(dWt + e, acc[j]);
    }
}

}  // namespace
}  // namespace mgar

using namespace mgar;

extern "C" size_t mgar_sa_mlp_train_workspace_bytes(int M, int nsample, int n_layers, const int *channels_host) {
    TrainSpec s;
    if (make_spec(s, M, nsample, n_layers, channels_host)) return 0;
    return s.total;
}

// Same, given the pooling method: the hot-path shape with max-pool runs the specialised kernels of
// sa_mlp_train_fast.cu, which keep no (positions x channels) activation (27 MB instead of 860 MB per JRDB frame).
extern "C" size_t mgar_sa_mlp_train_workspace_bytes_for(int M, int nsample, int n_layers, const int *channels_host, int pool_avg) {
    if (trfast::applicable(M, nsample, n_layers, channels_host, pool_avg)) return trfast::workspace_bytes(M, channels_host[0]);
    return mgar_sa_mlp_train_workspace_bytes(M, nsample, n_layers, channels_host);
}

static int tr_grid(long long work_items, int per_cta) {
    long long g = (work_items + per_cta - 1) / per_cta;
    const long long cap = (long long)kNumSMs * 8;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

extern "C" int mgar_sa_mlp_fwd_train(int B, int M, int nsample, int C_in, const float *xyz, const int *xyz_batch_cnt,
                                     const float *new_xyz, const int *new_xyz_batch_cnt, int uniform_n, int uniform_m,
                                     const float *features, const int *idx, const unsigned char *empty_mask,
                                     int n_layers, const int *channels_host, const float *weights,
                                     const float *gamma_beta, float eps, float *out, int ld_out, int pool_avg,
                                     float *batch_stats, void *workspace, size_t ws_bytes, void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0 && nsample >= 1 && C_in >= 0, "sa_mlp_train: bad size");
    TrainSpec s;
    MGAR_REQUIRE(make_spec(s, M, nsample, n_layers, channels_host) == 0, "sa_mlp_train: bad layer spec (1..%d layers)", kTrMaxLayers);
    if (M == 0) return MGAR_OK;
    MGAR_REQUIRE(xyz && new_xyz && idx && weights && gamma_beta && out && workspace, "sa_mlp_train: null pointer");
    MGAR_REQUIRE(features || C_in == 0, "sa_mlp_train: features missing");
    MGAR_REQUIRE(s.ch[0] == 3 + C_in, "sa_mlp_train: channels[0]=%d but 3+C_in=%d", s.ch[0], 3 + C_in);
    const bool uniform = uniform_n > 0;
    MGAR_REQUIRE(uniform || (xyz_batch_cnt && new_xyz_batch_cnt), "sa_mlp_train: batch counts missing");
    cudaStream_t st = (cudaStream_t)stream;
    if (trfast::applicable(M, nsample, n_layers, channels_host, pool_avg))
        return trfast::forward(B, M, C_in, xyz, xyz_batch_cnt, new_xyz, new_xyz_batch_cnt, uniform_n, uniform_m, features, idx,
                               empty_mask, weights, gamma_beta, eps, out, ld_out, batch_stats, workspace, ws_bytes, st);
    if (ws_bytes < s.total) { set_error("sa_mlp_train: workspace %zu < %zu bytes", ws_bytes, s.total); return MGAR_ENOMEM; }
    PrefixWs ws;
    if (!uniform) {   // ragged batches: the prefix arrays sit behind the training workspace (see the header)
        const size_t off = (s.total + 15) & ~(size_t)15;
        if (int rc = ws.bind(ws_bytes > off ? static_cast<char *>(workspace) + off : nullptr, ws_bytes > off ? ws_bytes - off : 0, B, "sa_mlp_fwd_train")) return rc;
        if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    }
    char *W = static_cast<char *>(workspace);
    float *x0 = reinterpret_cast<float *>(W + s.x0);
    tr_group_kernel<<<tr_grid(M, kTrThreads / 32), kTrThreads, 0, st>>>(B, M, nsample, C_in, xyz, new_xyz, features, ws.b, ws.a,
                                                                        uniform ? 1 : 0, uniform_n, uniform_m, idx, empty_mask, x0);
    count_launch();
    MGAR_CUDA(cudaMemsetAsync(W + s.stat64, 0, s.stat32 - s.stat64, st));
    const float *prev = x0, *prev_ss = nullptr;
    for (int l = 0; l < s.L; l++) {
        const int Ci = s.ch[l], Co = s.ch[l + 1];
        float *Z = reinterpret_cast<float *>(W + s.z[l]);
        int TP = 64;
        while (TP > 1 && (size_t)TP * (Ci + 1) * 4 > 96 * 1024) TP >>= 1;
        const size_t sm = (size_t)TP * (Ci + 1) * 4;
        if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(tr_linear_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        tr_linear_kernel<<<tr_grid(s.P, TP), kTrThreads, sm, st>>>(s.P, Ci, Co, TP, prev, prev_ss, prev_ss ? prev_ss + Ci : nullptr,
                                                                  weights + s.w_off[l], Z);
        count_launch();
        double *sum = reinterpret_cast<double *>(W + s.stat64) + s.c_off[l];
        float *mi = reinterpret_cast<float *>(W + s.stat32) + s.c_off[l];
        float *ss = reinterpret_cast<float *>(W + s.ss32) + s.c_off[l];
        const int rows = 2048;
        tr_colstats_kernel<<<(int)((s.P + rows - 1) / rows), kTrThreads, 0, st>>>(s.P, Co, rows, Z, sum, sum + Co);
        count_launch();
        tr_finalize_kernel<<<div_up(Co, 128), 128, 0, st>>>(s.P, Co, eps, sum, sum + Co, gamma_beta + s.g_off[l],
                                                             gamma_beta + s.g_off[l] + Co, mi, ss,
                                                             batch_stats ? batch_stats + s.c_off[l] : nullptr);
        count_launch();
        prev = Z; prev_ss = ss;
    }
    const int CL = s.ch[s.L];
    tr_pool_kernel<<<div_up((long long)M * CL, kTrThreads), kTrThreads, 0, st>>>(M, nsample, CL, prev, prev_ss, prev_ss + CL, out, ld_out,
                                                                                 reinterpret_cast<int *>(W + s.argmax), pool_avg);
    count_launch();
    return check_launch("sa_mlp_fwd_train");
}

extern "C" int mgar_sa_mlp_bwd(int B, int M, int nsample, int C_in, const int *xyz_batch_cnt, const int *new_xyz_batch_cnt,
                               int uniform_n, int uniform_m, const int *idx, const unsigned char *empty_mask, int n_layers,
                               const int *channels_host, const float *weights, const float *gamma_beta,
                               const float *grad_out, int ld_grad, int pool_avg, void *workspace, size_t ws_bytes,
                               float *grad_weights, float *grad_gamma_beta, float *grad_features, void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0 && nsample >= 1 && C_in >= 0, "sa_mlp_bwd: bad size");
    TrainSpec s;
    MGAR_REQUIRE(make_spec(s, M, nsample, n_layers, channels_host) == 0, "sa_mlp_bwd: bad layer spec");
    MGAR_REQUIRE(weights && gamma_beta && grad_out && workspace && grad_weights && grad_gamma_beta && idx, "sa_mlp_bwd: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (trfast::applicable(M, nsample, n_layers, channels_host, pool_avg)) {
        MGAR_REQUIRE(uniform_n > 0 || !grad_features || (xyz_batch_cnt && new_xyz_batch_cnt), "sa_mlp_bwd: batch counts missing");
        return trfast::backward(B, M, C_in, xyz_batch_cnt, new_xyz_batch_cnt, uniform_n, uniform_m, idx, empty_mask, weights,
                                gamma_beta, grad_out, ld_grad, workspace, ws_bytes, grad_weights, grad_gamma_beta,
                                grad_features, st);
    }
    if (ws_bytes < s.total) { set_error("sa_mlp_bwd: workspace %zu < %zu bytes", ws_bytes, s.total); return MGAR_ENOMEM; }
    int n_w = 0;
    for (int l = 0; l < s.L; l++) n_w += s.ch[l] * s.ch[l + 1];
    MGAR_CUDA(cudaMemsetAsync(grad_weights, 0, (size_t)n_w * 4, st));
    if (M == 0) {
        MGAR_CUDA(cudaMemsetAsync(grad_gamma_beta, 0, (size_t)(s.g_off[s.L - 1] + 2 * s.ch[s.L]) * 4, st));
        return MGAR_OK;
    }
    const bool uniform = uniform_n > 0;
    MGAR_REQUIRE(uniform || (xyz_batch_cnt && new_xyz_batch_cnt), "sa_mlp_bwd: batch counts missing");
    PrefixWs ws;
    if (!uniform && grad_features && C_in > 0) {
        const size_t off = (s.total + 15) & ~(size_t)15;
        if (int rc = ws.bind(ws_bytes > off ? static_cast<char *>(workspace) + off : nullptr, ws_bytes > off ? ws_bytes - off : 0, B, "sa_mlp_bwd")) return rc;
        if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    }
    char *W = static_cast<char *>(workspace);
    double *bs = reinterpret_cast<double *>(W + s.bstat64);
    float *coef = reinterpret_cast<float *>(W + s.coef);
    GradSrc gs;
    gs.buf = nullptr; gs.grad_out = grad_out; gs.arg = reinterpret_cast<const int *>(W + s.argmax);
    gs.ld = ld_grad; gs.K = nsample; gs.pool_avg = pool_avg;
    for (int l = s.L - 1; l >= 0; l--) {
        const int Ci = s.ch[l], Co = s.ch[l + 1];
        const float *Z = reinterpret_cast<const float *>(W + s.z[l]);
        const float *mi = reinterpret_cast<const float *>(W + s.stat32) + s.c_off[l];
        const float *ss = reinterpret_cast<const float *>(W + s.ss32) + s.c_off[l];
        MGAR_CUDA(cudaMemsetAsync(bs, 0, (size_t)2 * Co * 8, st));
        const int rows = 2048;
        tr_bwd_stats_kernel<<<(int)((s.P + rows - 1) / rows), kTrThreads, 0, st>>>(s.P, Co, rows, Z, mi, ss, gs, bs, bs + Co);
        count_launch();
        tr_bwd_coef_kernel<<<div_up(Co, 128), 128, 0, st>>>(s.P, Co, bs, bs + Co, gamma_beta + s.g_off[l], mi, coef,
                                                             grad_gamma_beta + s.g_off[l], grad_gamma_beta + s.g_off[l] + Co);
        count_launch();
        const float *Aprev = l == 0 ? reinterpret_cast<const float *>(W + s.x0) : reinterpret_cast<const float *>(W + s.z[l - 1]);
        const float *ss_prev = l == 0 ? nullptr : reinterpret_cast<const float *>(W + s.ss32) + s.c_off[l - 1];
        float *g_prev = l == 0 ? nullptr : reinterpret_cast<float *>(W + s.gbuf[l & 1]);
        ScatterArgs sa;
        sa.grad_features = (l == 0 && C_in > 0) ? grad_features : nullptr;
        sa.C_in = C_in; sa.B = B; sa.uniform = uniform ? 1 : 0; sa.n_u = uniform_n; sa.m_u = uniform_m;
        sa.pre_xyz = ws.b; sa.pre_new = ws.a; sa.idx = idx; sa.empty_mask = empty_mask;
        size_t sm = ((size_t)kTrTP * (Co + 1) + (size_t)kTrTP * (Ci + 1)) * 4;
        const size_t smw = (size_t)Ci * (Co + 1) * 4;
        const int w_in_smem = (sm + smw <= 160 * 1024) ? 1 : 0;
        if (w_in_smem) sm += smw;
        MGAR_REQUIRE(sm <= 200 * 1024, "sa_mlp_bwd: layer %d needs %zu bytes of shared memory", l, sm);
        if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(tr_bwd_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        const long long pairs = (long long)Ci * Co;
        dim3 grid(tr_grid(s.P, kTrTP * 16) > kNumSMs * 2 ? kNumSMs * 2 : tr_grid(s.P, kTrTP * 16),
                  (unsigned)((pairs + (long long)kTrThreads * kTrAcc - 1) / ((long long)kTrThreads * kTrAcc)));
        tr_bwd_layer_kernel<<<grid, kTrThreads, sm, st>>>(s.P, Ci, Co, Z, mi, ss, coef, gs, Aprev, ss_prev, ss_prev ? ss_prev + Ci : nullptr,
                                                          weights + s.w_off[l], w_in_smem, grad_weights + s.w_off[l], g_prev, sa);
        count_launch();
        gs.buf = g_prev;
    }
    return check_launch("sa_mlp_bwd");
}
