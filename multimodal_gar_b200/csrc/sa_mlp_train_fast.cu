// Training step of the hot-path SA shared MLP, shape [3+C_in, 64, 128], nsample 32, max-pool: BatchNorm with batch
// statistics forward + the whole backward WITHOUT materialising any (positions x channels) activation.
//
// Reference: StackSAModuleMSG.forward in train() (pointnet2_stack/pointnet2_modules.py:78-112) and the backward
// torch autograd derives for it (Conv2d 1x1 / BatchNorm2d / ReLU / max_pool2d + GroupingOperation.backward,
// group_points_gpu.cu:15-44).  The generic kernels in sa_mlp_train.cu keep every pre-activation (860 MB per
// JRDB frame) and run two dense (positions x 64 x 128) contractions in the backward.  Three identities remove
// both for this shape (P = M*32 positions, x0 = grouped input rows, h1 = relu(bn1(x0 W1)), z2 = h1 W2):
//
//  (1) BatchNorm's affine is monotone, rounding is monotone: max_k relu(a z_k + b) = relu(a zext + b) with
//      zext = max_k z_k (a >= 0) or min_k z_k (a < 0), and sign(a) = sign(gamma2) is known before the statistics.
//      The forward keeps that extreme and its sample per (centroid, channel) plus the column sums of z2, z2^2 -
//      z2 itself is never written.  (Tie rule: first k attaining the
//      extreme z; equal to torch's "first k attaining the max y" except when two DIFFERENT z round to the same
//      y > 0, or gamma == 0 exactly.)
//  (2) Layer-1 statistics are a quadratic form: sum_p z1 = W1 . X, sum_p z1^2 = W1^T G0 W1 with X = sum_p x0,
//      G0 = x0^T x0 (C0 x C0, fp64).
//  (3) The gradient entering BatchNorm 2 is SPARSE (one non-zero per (centroid, channel): the arg-max sample), and
//      g_z2 = coefA (g_y - s1/P - xhat s2/P) is that sparse term plus a term LINEAR in z2 = h1 W2:
//        dW2^T = A - u H^T ... - e (G W2 - mean2 H)      A = sum over arg-max rows, H = sum_p h1, G = h1^T h1 (64x64)
//        g_h1  = sparse rows . W2^T - h1 Q + cvec        Q = W2 diag(e) W2^T (64x64)
//      so the backward's dense work is one (P x 64 x 64) product plus the Gram matrix.  The same identity gives
//      layer 1: dW1, dgamma1, dbeta1 from D = g_y1^T x0 (64 x C0), s1, X, G0, and the feature gradient
//      t[p] - x0[p] . q + c with t[p] = sum_c alpha[c] g_y1[p, c].
//
// fp32 FFMA contractions on CUDA cores (register-tiled, operands in shared memory), fp64 for every batch reduction.
// Test: tests/test_train_gpu.py (outputs / running statistics 1e-5, gradients 1e-4 against torch autograd).
#include "sa_mlp_train_fast.cuh"

#include <stdlib.h>
#include <string.h>

#include "sa_mlp_tc.cuh"
#include "stack_prefix.cuh"

namespace mgar {
namespace trfast {

namespace {

constexpr int kC1 = 64, kC2 = 128, kK = 32, kTP = 128, kThreads = 256, kLdH = 68;
constexpr int kStatEnt = kC1 * kC1 + kC1 * kC2 + kC1;  // G, A^T, H
constexpr int kMaxGrid = kNumSMs * 2;

// fp64 block (indices in doubles)
constexpr int dX = 0, dG0 = 8, dSum2 = 72, dSq2 = 200, dFwdEnd = 328;
constexpr int dS1 = 328, dS2 = 456, dD = 584, dS11 = 1096, dRed = 1160, dTotal = dRed + kStatEnt;
// fp32 block (indices in floats)
constexpr int fMean1 = 0, fInv1 = 64, fSc1 = 128, fSh1 = 192, fAlpha = 768, fMean2 = 1088, fInv2 = 1216,
              fSc2 = 1344, fSh2 = 1472, fQ = 1600, fCvec = 5696, fE = 5760, fU = 5888, fQf = 6016, fCf = 6056,
              fSgn2 = 6064, fZero2 = 6192, fL1 = 6320 /* 8 x 64 folded W1 | 64 shifts | wmax[8], shmax, pad */, fTotal = 6320 + 9 * 64 + 16;

struct Layout {
    size_t x0, zmax, gs, arg, t, lst, lgs, lcnt, part, dbl, flt, w2t, total;
};

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

Layout make_layout(int M, int C0) {
    Layout l;
    const size_t P = (size_t)M * kK;
    size_t off = 0;
    l.x0 = off; off = align_up(off + P * C0 * 4);
    l.zmax = off; off = align_up(off + (size_t)M * kC2 * 4);
    l.gs = off;   off = align_up(off + (size_t)M * kC2 * 4);
    l.arg = off; off = align_up(off + (size_t)M * kC2 * 4);
    l.t = off; off = align_up(off + P * (C0 - 3) * 4);
    l.lst = off; off = align_up(off + (size_t)M * kC2 * 4);
    l.lgs = off; off = align_up(off + (size_t)M * kC2 * 4);
    l.lcnt = off; off = align_up(off + (size_t)M * 4);
    l.part = off; off = align_up(off + (size_t)kMaxGrid * kStatEnt * 8);
    l.dbl = off; off = align_up(off + (size_t)dTotal * 8);
    l.flt = off; off = align_up(off + (size_t)fTotal * 4);
    l.w2t = off; off = align_up(off + (size_t)kC1 * kC2 * 4);
    l.total = off;
    return l;
}

__device__ __forceinline__ int find_batch(const int *__restrict__ pre, int B, int v) {
    int lo = 0, hi = B;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(pre + mid) <= v) lo = mid; else hi = mid;
    }
    return lo;
}

// Packed fp32 pairs: Blackwell's FFMA2 (fma.rn.f32x2) does two IEEE fp32 FMAs per lane per issue slot - the scalar
// FFMA issues every other cycle per scheduler, so the register-tiled contractions below run on pairs.
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t pack2(float lo, float hi) {
    f2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f2_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void ffma2(f2_t &d, f2_t a, f2_t b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- forward ---------------------------------------------------------------------------------------------------
// x0[p, :] = (xyz[src] - centre | features[src]) or zeros for an empty ball (pointnet2_utils.py:146-150);
// X = sum_p x0, G0 = x0^T x0 in fp64.  Warp per centroid, lane = sample.
template <int C0>
__global__ void __launch_bounds__(kThreads)
f_group_kernel(int B, int M, const float *__restrict__ xyz, const float *__restrict__ new_xyz,
               const float *__restrict__ features, const int *__restrict__ pre_xyz, const int *__restrict__ pre_new,
               int uniform, int n_u, int m_u, const int *__restrict__ idx, const unsigned char *__restrict__ empty_mask,
               float *__restrict__ x0, double *__restrict__ dbl) {
    constexpr int C_in = C0 - 3, NT = C0 * (C0 + 1) / 2;
    __shared__ double red[kThreads / 32][C0 + NT];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double sx[C0], sg[NT];
#pragma unroll
    for (int i = 0; i < C0; i++) sx[i] = 0.0;
#pragma unroll
    for (int i = 0; i < NT; i++) sg[i] = 0.0;
    for (int m = blockIdx.x * (kThreads / 32) + warp; m < M; m += gridDim.x * (kThreads / 32)) {
        const bool empty = (__ldg(idx + (size_t)m * kK) < 0) || (empty_mask && empty_mask[m]);
        float row[C0];
#pragma unroll
        for (int i = 0; i < C0; i++) row[i] = 0.f;
        if (!empty) {
            const long long start = uniform ? (long long)(m / m_u) * n_u : __ldg(pre_xyz + find_batch(pre_new, B, m));
            const long long p = start + __ldg(idx + (size_t)m * kK + lane);
#pragma unroll
            for (int i = 0; i < 3; i++) row[i] = __fsub_rn(__ldg(xyz + p * 3 + i), __ldg(new_xyz + (size_t)m * 3 + i));
#pragma unroll
            for (int c = 0; c < C_in; c++) row[3 + c] = __ldg(features + p * C_in + c);
        }
        float *dst = x0 + ((size_t)m * kK + lane) * C0;
        if (C0 == 4) {
            *reinterpret_cast<float4 *>(dst) = make_float4(row[0], row[1], row[2], row[3]);
        } else {
#pragma unroll
            for (int i = 0; i < C0; i++) dst[i] = row[i];
        }
        int tt = 0;
#pragma unroll
        for (int i = 0; i < C0; i++) {
            sx[i] += (double)row[i];
#pragma unroll
            for (int j = i; j < C0; j++) sg[tt++] += (double)row[i] * (double)row[j];
        }
    }
#pragma unroll
    for (int i = 0; i < C0; i++) { const double v = warp_sum(sx[i]); if (lane == 0) red[warp][i] = v; }
#pragma unroll
    for (int i = 0; i < NT; i++) { const double v = warp_sum(sg[i]); if (lane == 0) red[warp][C0 + i] = v; }
    __syncthreads();
    if (threadIdx.x < C0 + NT) {
        double v = 0.0;
        for (int w = 0; w < kThreads / 32; w++) v += red[w][threadIdx.x];
        if (threadIdx.x < C0) {
            atomicAdd(dbl + dX + threadIdx.x, v);
        } else {
            int tt = threadIdx.x - C0, i = 0;
            while (tt >= C0 - i) { tt -= C0 - i; i++; }
            atomicAdd(dbl + dG0 + i * 8 + (i + tt), v);
        }
    }
}

__device__ __forceinline__ double g0_sym(const double *__restrict__ dbl, int i, int j) {
    return i <= j ? dbl[dG0 + i * 8 + j] : dbl[dG0 + j * 8 + i];
}

// BatchNorm 1 from (X, G0): mean, biased variance, scale / shift, alpha[f][c] = scale1[c] W1[3+f][c]
template <int C0>
__global__ void f_bn1_kernel(long long P, float eps, const float *__restrict__ Wt1, const float *__restrict__ gamma_beta,
                             const double *__restrict__ dbl, float *__restrict__ flt, float *__restrict__ batch_stats) {
    const int c = threadIdx.x;
    if (c >= kC1) return;
    double w[C0], mean = 0.0, ez2 = 0.0;
#pragma unroll
    for (int i = 0; i < C0; i++) { w[i] = (double)Wt1[i * kC1 + c]; mean += w[i] * dbl[dX + i]; }
    mean /= (double)P;
#pragma unroll
    for (int i = 0; i < C0; i++)
#pragma unroll
        for (int j = 0; j < C0; j++) ez2 += w[i] * w[j] * g0_sym(dbl, i, j);
    double var = ez2 / (double)P - mean * mean;
    if (var < 0.0) var = 0.0;
    const double inv = 1.0 / sqrt(var + (double)eps);
    const double sc = (double)gamma_beta[c] * inv;
    flt[fMean1 + c] = (float)mean;
    flt[fInv1 + c] = (float)inv;
    flt[fSc1 + c] = (float)sc;
    flt[fSh1 + c] = (float)((double)gamma_beta[kC1 + c] - mean * sc);
#pragma unroll
    for (int f = 0; f < C0 - 3; f++) flt[fAlpha + f * kC1 + c] = (float)(sc * w[3 + f]);
    if (batch_stats) { batch_stats[c] = (float)mean; batch_stats[kC1 + c] = (float)var; }
    for (int k = c; k < kC2; k += kC1) {      // sign(gamma2) and a zero shift: the tensor-core statistics pass
        flt[fSgn2 + k] = gamma_beta[2 * kC1 + k] < 0.f ? -1.f : 1.f;
        flt[fZero2 + k] = 0.f;
    }
    // ... and its layer 1 (TcLayer1 / TcBound of sa_mlp_tc.cuh) on the device: W1 with BatchNorm 1's batch scale folded in,
    // the shift, and the per-input-column bound max_c |w[k][c]|, max_c |shift[c]|
    __shared__ float s_red[2][9];
    const float scf = (float)sc, shf = (float)((double)gamma_beta[kC1 + c] - mean * sc);
    float wm[9];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float v = k < C0 ? __fmul_rn(Wt1[k * kC1 + c], scf) : 0.f;
        flt[fL1 + k * kC1 + c] = v;
        wm[k] = fabsf(v);
    }
    flt[fL1 + 8 * kC1 + c] = shf;
    wm[8] = fabsf(shf);
#pragma unroll
    for (int k = 0; k < 9; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wm[k] = fmaxf(wm[k], __shfl_xor_sync(0xffffffffu, wm[k], o));
        if ((c & 31) == 0) s_red[c >> 5][k] = wm[k];
    }
    __syncthreads();
    if (c < 9) flt[fL1 + 9 * kC1 + c] = fmaxf(s_red[0][c], s_red[1][c]);
}

// h1 rows of one 128-position tile: thread (c1 = t & 63, q = t >> 6) writes channel c1 of the 32 rows of centroid q.
// Same arithmetic as the generic kernels (tr_linear_kernel + its consumer): z1 = FMA chain over the inputs from 0,
// h1 = relu(fma(z1, scale1, shift1)) - so both native paths take the same ReLU decisions.
struct L1Params { float sc, sh; };
template <int C0>
__device__ __forceinline__ void load_l1(const float *__restrict__ Wt1, const float *__restrict__ flt, int c1, float (&w1)[C0], L1Params &p) {
#pragma unroll
    for (int i = 0; i < C0; i++) w1[i] = __ldg(Wt1 + i * kC1 + c1);
    p.sc = __ldg(flt + fSc1 + c1);
    p.sh = __ldg(flt + fSh1 + c1);
}
template <int C0>
__device__ __forceinline__ void h1_tile(const float *__restrict__ sX, float *__restrict__ sH, const float (&w1)[C0],
                                        L1Params lp, int c1, int q, bool valid) {
#pragma unroll 8
    for (int r = 0; r < kK; r++) {
        const int pos = q * kK + r;
        float z = 0.f;
#pragma unroll
        for (int i = 0; i < C0; i++) z = __fmaf_rn(sX[pos * C0 + i], w1[i], z);
        sH[pos * kLdH + c1] = valid ? fmaxf(__fmaf_rn(z, lp.sc, lp.sh), 0.f) : 0.f;
    }
}

template <int C0>
__device__ __forceinline__ void load_x_tile(const float *__restrict__ x0, float *__restrict__ sX, long long p0, long long P) {
    const long long base = p0 * C0, lim = P * C0;
    for (int e = threadIdx.x; e < kTP * C0; e += kThreads) sX[e] = (base + e < lim) ? __ldg(x0 + base + e) : 0.f;
}

// z2 = h1 W2 per 128-position tile (never stored): column sums of z2 and z2^2 (fp64 slots in shared memory), and
// per (centroid, channel) the extreme of z2 over the 32 samples that survives the max-pool, with the first sample
// attaining it: the max where gamma2 >= 0, the min where gamma2 < 0.  The sign is folded into the W2 columns held in
// shared memory (exact), so the kernel tracks max(sgn z2) only and zmax holds sgn * z2.
// Thread tile 8 positions x 8 channels; warp = one centroid x 64 channels.
template <int C0>
__global__ void __launch_bounds__(kThreads, 2)
f_layer2_kernel(long long P, int M, const float *__restrict__ x0, const float *__restrict__ flt,
                const float *__restrict__ Wt1, const float *__restrict__ Wt2, const float *__restrict__ gamma2,
                float *__restrict__ zmax, int *__restrict__ arg, double *__restrict__ dbl) {
    extern __shared__ __align__(16) unsigned char tf_sm[];
    float *sW2 = reinterpret_cast<float *>(tf_sm);            // [64][128]
    float *sH = sW2 + kC1 * kC2;                              // [128][68]
    float *sX = sH + kTP * kLdH;                              // [128][C0] (room for 8)
    double *sD = reinterpret_cast<double *>(sX + kTP * 8);    // [8 warps][8 txl][16]
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int e = t; e < kC1 * kC2; e += kThreads) sW2[e] = __ldg(gamma2 + (e & (kC2 - 1))) < 0.f ? -__ldg(Wt2 + e) : __ldg(Wt2 + e);
    for (int e = t; e < 8 * 8 * 16; e += kThreads) sD[e] = 0.0;
    const int c1 = t & 63, q = t >> 6;
    float w1[C0];
    L1Params lp;
    load_l1<C0>(Wt1, flt, c1, w1, lp);
    const int wy = warp >> 1, wx = warp & 1, tyl = lane >> 3, txl = lane & 7;
    const long long ntiles = (P + kTP - 1) / kTP;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        __syncthreads();
        load_x_tile<C0>(x0, sX, tile * kTP, P);
        __syncthreads();
        h1_tile<C0>(sX, sH, w1, lp, c1, q, tile * 4 + q < M);
        __syncthreads();
        f2_t acc2[8][4];
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc2[r][j] = 0ull;
        const float *aBase = sH + (wy * kK + tyl) * kLdH;   // samples k = 4 r + tyl: the 4 lane groups read adjacent rows (distinct banks)
        const float *bBase = sW2 + wx * 64 + txl * 4;
#pragma unroll 1
        for (int k4 = 0; k4 < kC1 / 4; k4++) {
            float4 a[8];
#pragma unroll
            for (int r = 0; r < 8; r++) a[r] = *reinterpret_cast<const float4 *>(aBase + r * 4 * kLdH + k4 * 4);
#pragma unroll
            for (int kk = 0; kk < 4; kk++) {
                const float4 b0 = *reinterpret_cast<const float4 *>(bBase + (k4 * 4 + kk) * kC2);
                const float4 b1 = *reinterpret_cast<const float4 *>(bBase + (k4 * 4 + kk) * kC2 + 32);
                const f2_t b01 = pack2(b0.x, b0.y), b23 = pack2(b0.z, b0.w), b45 = pack2(b1.x, b1.y), b67 = pack2(b1.z, b1.w);
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    const float av = kk == 0 ? a[r].x : kk == 1 ? a[r].y : kk == 2 ? a[r].z : a[r].w;
                    const f2_t aa = pack2(av, av);
                    ffma2(acc2[r][0], aa, b01);
                    ffma2(acc2[r][1], aa, b23);
                    ffma2(acc2[r][2], aa, b45);
                    ffma2(acc2[r][3], aa, b67);
                }
            }
        }
        // epilogue: statistics over this thread's 8 samples, then over the 4 lanes (tyl) that share the channels
        float mx[8], s[8], sq[8];
        int ax[8];
        double *slot = sD + (warp * 8 + txl) * 16;
        const f2_t one2 = pack2(1.f, 1.f);
#pragma unroll
        for (int jp = 0; jp < 4; jp++) {
            f2_t s2 = 0ull, q2 = 0ull;
#pragma unroll
            for (int r = 0; r < 8; r++) { ffma2(s2, acc2[r][jp], one2); ffma2(q2, acc2[r][jp], acc2[r][jp]); }
            unpack2(s2, s[2 * jp], s[2 * jp + 1]);
            unpack2(q2, sq[2 * jp], sq[2 * jp + 1]);
        }
#pragma unroll
        for (int j = 0; j < 8; j++) {
            float v0, v1;
            unpack2(acc2[0][j >> 1], v0, v1);
            mx[j] = (j & 1) ? v1 : v0; ax[j] = 0;
#pragma unroll
            for (int r = 1; r < 8; r++) {
                unpack2(acc2[r][j >> 1], v0, v1);
                const float v = (j & 1) ? v1 : v0;
                if (v > mx[j]) { mx[j] = v; ax[j] = r; }
            }
            ax[j] = ax[j] * 4 + tyl;
#pragma unroll
            for (int o = 8; o <= 16; o <<= 1) {
                s[j] += __shfl_xor_sync(0xffffffffu, s[j], o);
                sq[j] += __shfl_xor_sync(0xffffffffu, sq[j], o);
                const float omx = __shfl_xor_sync(0xffffffffu, mx[j], o);
                const int oax = __shfl_xor_sync(0xffffffffu, ax[j], o);
                if (omx > mx[j] || (omx == mx[j] && oax < ax[j])) { mx[j] = omx; ax[j] = oax; }
            }
            if (tyl == 0) { slot[j] += (double)s[j]; slot[8 + j] += (double)sq[j]; }
        }
        const long long m = tile * 4 + wy;
        if (tyl == 0 && m < M) {
            const size_t o = (size_t)m * kC2 + wx * 64 + txl * 4;
            *reinterpret_cast<float4 *>(zmax + o) = make_float4(mx[0], mx[1], mx[2], mx[3]);
            *reinterpret_cast<float4 *>(zmax + o + 32) = make_float4(mx[4], mx[5], mx[6], mx[7]);
            *reinterpret_cast<int4 *>(arg + o) = make_int4(ax[0], ax[1], ax[2], ax[3]);
            *reinterpret_cast<int4 *>(arg + o + 32) = make_int4(ax[4], ax[5], ax[6], ax[7]);
        }
    }
    __syncthreads();
    {   // t -> (wx, txl, jj): channel and sum / sum-of-squares selector
        const int wxx = t >> 7, rest = t & 127, txx = rest >> 4, jj = rest & 15, j = jj & 7;
        const int c = wxx * 64 + txx * 4 + (j & 3) + (j >> 2) * 32;
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < 4; w++) v += sD[((w * 2 + wxx) * 8 + txx) * 16 + jj];
        if (jj < 8 && __ldg(gamma2 + c) < 0.f) v = -v;      // the columns were negated: sum z2 = -sum(sgn z2)
        atomicAdd(dbl + (jj >= 8 ? dSq2 : dSum2) + c, v);
    }
}

// Entries the tensor-core statistics pass flagged (a runner-up sample within the 3xFP16 split's error of the
// extreme): the 32 samples are re-evaluated in exact fp32 with f_layer2_kernel's arithmetic (h1 chain, then z2 as an
// FMA chain over the 64 inputs in ascending order) and the extreme / first sample are rewritten.  Warp per entry,
// lane = sample; a few entries per thousand.
template <int C0>
__global__ void __launch_bounds__(kThreads)
f_fix_ties_kernel(long long n, const float *__restrict__ x0, const float *__restrict__ flt, const float *__restrict__ Wt1,
                  const float *__restrict__ Wt2, const float *__restrict__ gamma2, float *__restrict__ zmax,
                  int *__restrict__ arg) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * kThreads + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * kThreads) >> 5;
    for (long long e0 = warp * 32; e0 < n; e0 += nwarps * 32) {
        const long long e = e0 + lane;
        const int a = e < n ? arg[e] : 0;
        unsigned todo = __ballot_sync(0xffffffffu, (a & 0x100) != 0);
        while (todo) {
            const int b = __ffs(todo) - 1;
            todo &= todo - 1;
            const long long ee = e0 + b, m = ee >> 7;
            const int c = (int)(ee & (kC2 - 1));
            float x[C0];
#pragma unroll
            for (int i = 0; i < C0; i++) x[i] = __ldg(x0 + (m * kK + lane) * C0 + i);
            float z2 = 0.f;
            for (int j = 0; j < kC1; j++) {
                float z1 = 0.f;
#pragma unroll
                for (int i = 0; i < C0; i++) z1 = __fmaf_rn(x[i], __ldg(Wt1 + i * kC1 + j), z1);
                const float h = fmaxf(__fmaf_rn(z1, __ldg(flt + fSc1 + j), __ldg(flt + fSh1 + j)), 0.f);
                z2 = __fmaf_rn(h, __ldg(Wt2 + j * kC2 + c), z2);
            }
            float v = __ldg(gamma2 + c) < 0.f ? -z2 : z2;
            int k = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, v, o);
                const int ok = __shfl_xor_sync(0xffffffffu, k, o);
                if (ov > v || (ov == v && ok < k)) { v = ov; k = ok; }
            }
            if (lane == 0) { zmax[ee] = v; arg[ee] = k; }
        }
        __syncwarp();
    }
}

__global__ void f_bn2_kernel(long long P, float eps, const float *__restrict__ gamma_beta, const double *__restrict__ dbl,
                             float *__restrict__ flt, float *__restrict__ batch_stats) {
    const int c = threadIdx.x;
    if (c >= kC2) return;
    const double mean = dbl[dSum2 + c] / (double)P;
    double var = dbl[dSq2 + c] / (double)P - mean * mean;
    if (var < 0.0) var = 0.0;
    const double inv = 1.0 / sqrt(var + (double)eps);
    const double sc = (double)gamma_beta[2 * kC1 + c] * inv;
    flt[fMean2 + c] = (float)mean;
    flt[fInv2 + c] = (float)inv;
    flt[fSc2 + c] = (float)sc;
    flt[fSh2 + c] = (float)((double)gamma_beta[2 * kC1 + kC2 + c] - mean * sc);
    if (batch_stats) { batch_stats[2 * kC1 + c] = (float)mean; batch_stats[2 * kC1 + kC2 + c] = (float)var; }
}

// out = relu(scale2 * zext + shift2); zmax <- zext (the sign folded by the layer-2 pass is undone here).
// After the tensor-core pass (x0 != nullptr) an entry whose BatchNorm output is within rounding of zero gets its z2
// re-evaluated in exact fp32 (f_layer2_kernel's arithmetic), so the ReLU decision is the CUDA-core path's.
template <int C0>
__global__ void __launch_bounds__(kThreads)
f_pool_kernel(long long n, const float *__restrict__ flt, const float *__restrict__ gamma2, float *__restrict__ zmax,
              float *__restrict__ out, int ld_out, const float *__restrict__ x0, const int *__restrict__ arg,
              const float *__restrict__ Wt1, const float *__restrict__ Wt2) {
    const long long e = (long long)blockIdx.x * kThreads + threadIdx.x;
    if (e >= n) return;
    const int c = (int)(e & (kC2 - 1));
    const long long m = e >> 7;
    const float a = __ldg(flt + fSc2 + c), b = __ldg(flt + fSh2 + c);
    float z = __ldg(gamma2 + c) < 0.f ? -zmax[e] : zmax[e];
    float y = __fmaf_rn(z, a, b);
    if (x0 && fabsf(y) <= 1e-5f * (fabsf(z * a) + fabsf(b))) {
        const float *x = x0 + (m * kK + __ldg(arg + e)) * C0;
        float z2 = 0.f;
        for (int j = 0; j < kC1; j++) {
            float z1 = 0.f;
#pragma unroll
            for (int i = 0; i < C0; i++) z1 = __fmaf_rn(__ldg(x + i), __ldg(Wt1 + i * kC1 + j), z1);
            const float h = fmaxf(__fmaf_rn(z1, __ldg(flt + fSc1 + j), __ldg(flt + fSh1 + j)), 0.f);
            z2 = __fmaf_rn(h, __ldg(Wt2 + j * kC2 + c), z2);
        }
        z = z2;
        y = __fmaf_rn(z, a, b);
    }
    zmax[e] = z;
    out[m * ld_out + c] = fmaxf(y, 0.f);
}

// ---- backward --------------------------------------------------------------------------------------------------
// g_y at the arg-max samples, s1 = sum g_y, s2 = sum g_y xhat (fp64); gs = coefA * g_y.
// Warp per centroid, lane l owns channels l + 32 j.  The non-zero (channel, sample) entries of the centroid are also
// written as a list SORTED BY SAMPLE (stable in the channel; counting sort with match_any ranks): b_layer_kernel
// turns them into rows of g_h1 with register accumulation, one store per sample.  The list is padded to a multiple
// of 4 with zero-valued entries of sample 31.
__global__ void __launch_bounds__(kThreads)
b_gy_kernel(int M, const float *__restrict__ zsel, const int *__restrict__ arg, const float *__restrict__ flt,
            const float *__restrict__ grad_out, int ld, float *__restrict__ gs, unsigned *__restrict__ lst,
            float *__restrict__ lgs, int *__restrict__ lcnt, double *__restrict__ dbl) {
    __shared__ double acc_s[2][kC2];
    __shared__ int cnt_s[kThreads / 32][32], st_s[kThreads / 32][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < kC2) { acc_s[0][threadIdx.x] = 0.0; acc_s[1][threadIdx.x] = 0.0; }
    __syncthreads();
    float a[4], b[4], mean[4], inv[4];
    double s1[4], s2[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int c = lane + 32 * j;
        a[j] = __ldg(flt + fSc2 + c); b[j] = __ldg(flt + fSh2 + c);
        mean[j] = __ldg(flt + fMean2 + c); inv[j] = __ldg(flt + fInv2 + c);
        s1[j] = 0.0; s2[j] = 0.0;
    }
    const unsigned lt = (1u << lane) - 1u;
    for (long long m = (long long)blockIdx.x * (kThreads / 32) + warp; m < M; m += (long long)gridDim.x * (kThreads / 32)) {
        float gv[4];
        int kk[4], off[4];
        cnt_s[warp][lane] = 0;
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int c = lane + 32 * j;
            const float z = __ldg(zsel + m * kC2 + c);
            const float gy = __fmaf_rn(z, a[j], b[j]) > 0.f ? __ldg(grad_out + m * ld + c) : 0.f;
            gv[j] = a[j] * gy;
            kk[j] = __ldg(arg + m * kC2 + c);
            gs[m * kC2 + c] = gv[j];
            s1[j] += (double)gy;
            s2[j] += (double)gy * (double)((z - mean[j]) * inv[j]);
            const bool nz = gv[j] != 0.f;
            const unsigned peers = __match_any_sync(0xffffffffu, nz ? kk[j] : 64 + lane);
            const int base = cnt_s[warp][kk[j]];
            __syncwarp();
            if (nz && (peers & lt) == 0) cnt_s[warp][kk[j]] = base + __popc(peers);
            __syncwarp();
            off[j] = base + __popc(peers & lt);
        }
        // exclusive scan of the 32 sample buckets
        const int mine = cnt_s[warp][lane];
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        st_s[warp][lane] = incl - mine;
        const int n = __shfl_sync(0xffffffffu, incl, 31);
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (gv[j] != 0.f) {
                const int pos = st_s[warp][kk[j]] + off[j];
                lst[m * kC2 + pos] = ((unsigned)kk[j] << 8) | (unsigned)(lane + 32 * j);
                lgs[m * kC2 + pos] = gv[j];
            }
        }
        const int n4 = (n + 3) & ~3;
        if (lane < n4 - n) { lst[m * kC2 + n + lane] = 31u << 8; lgs[m * kC2 + n + lane] = 0.f; }
        if (lane == 0) lcnt[m] = n4;
        __syncwarp();
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
        atomicAdd(&acc_s[0][lane + 32 * j], s1[j]);
        atomicAdd(&acc_s[1][lane + 32 * j], s2[j]);
    }
    __syncthreads();
    if (threadIdx.x < kC2) {
        atomicAdd(dbl + dS1 + threadIdx.x, acc_s[0][threadIdx.x]);
        atomicAdd(dbl + dS2 + threadIdx.x, acc_s[1][threadIdx.x]);
    }
}

// e[c] = coefA (s2/P) invstd2, u[c] = coefA s1/P; Q = W2 diag(e) W2^T; cvec; W2 transposed; dgamma2, dbeta2
__global__ void __launch_bounds__(kThreads)
b_prep_kernel(long long P, const float *__restrict__ Wt2, const double *__restrict__ dbl, float *__restrict__ flt,
              float *__restrict__ W2T, float *__restrict__ gGB) {
    __shared__ float se[kC2], su[kC2];
    const int t = threadIdx.x;
    if (t < kC2) {
        const double a = (double)flt[fSc2 + t];
        se[t] = (float)(a * (dbl[dS2 + t] / (double)P) * (double)flt[fInv2 + t]);
        su[t] = (float)(a * (dbl[dS1 + t] / (double)P));
    }
    __syncthreads();
    if (blockIdx.x < 16) {
        const int e = blockIdx.x * kThreads + t, j = e >> 6, ci = e & 63;
        float s = 0.f;
        for (int c = 0; c < kC2; c++) s = __fmaf_rn(se[c] * __ldg(Wt2 + j * kC2 + c), __ldg(Wt2 + ci * kC2 + c), s);
        flt[fQ + e] = s;
    } else {
        if (t < kC1) {
            float s = 0.f;
            for (int c = 0; c < kC2; c++) s = __fmaf_rn(__ldg(Wt2 + t * kC2 + c), se[c] * flt[fMean2 + c] - su[c], s);
            flt[fCvec + t] = s;
        } else if (t < kC1 + kC2) {
            const int c = t - kC1;
            flt[fE + c] = se[c];
            flt[fU + c] = su[c];
            gGB[2 * kC1 + c] = (float)dbl[dS2 + c];
            gGB[2 * kC1 + kC2 + c] = (float)dbl[dS1 + c];
        }
        for (int e = t; e < kC1 * kC2; e += kThreads) W2T[e] = __ldg(Wt2 + (e & 63) * kC2 + (e >> 6));
    }
}

// G = h1^T h1, H = sum_p h1, A^T[c][ci] = sum over centroids gs[m, c] h1[arg-max row of (m, c)][ci], accumulated
// per CTA in fp32 registers and flushed to the CTA's private fp64 block every 16 tiles (no atomics except H).
template <int C0>
__global__ void __launch_bounds__(kThreads, 2)
b_h1stats_kernel(long long P, int M, const float *__restrict__ x0, const float *__restrict__ flt,
                 const float *__restrict__ Wt1, const float *__restrict__ gs, const int *__restrict__ arg, double *__restrict__ part) {
    extern __shared__ __align__(16) unsigned char tf_sm[];
    float *sH = reinterpret_cast<float *>(tf_sm);   // [128][68]
    float *sX = sH + kTP * kLdH;                    // [128][C0] (room for 8)
    float *sGs = sX + kTP * 8;                      // [4][128]
    int *sArg = reinterpret_cast<int *>(sGs + 4 * kC2);
    const int t = threadIdx.x;
    const int ci = t & 63, cg = t >> 6, gi = t >> 4, gj = t & 15;
    float w1[C0];
    L1Params lp;
    load_l1<C0>(Wt1, flt, ci, w1, lp);
    f2_t g2[4][2];
    float a[32], hs = 0.f;
#pragma unroll
    for (int u = 0; u < 4; u++) { g2[u][0] = 0ull; g2[u][1] = 0ull; }
#pragma unroll
    for (int i = 0; i < 32; i++) a[i] = 0.f;
    double *mine = part + (size_t)blockIdx.x * kStatEnt;
    auto flush = [&]() {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int v = 0; v < 2; v++) {
                float lo, hi;
                unpack2(g2[u][v], lo, hi);
                mine[(gi * 4 + u) * kC1 + gj * 4 + 2 * v] += (double)lo;
                mine[(gi * 4 + u) * kC1 + gj * 4 + 2 * v + 1] += (double)hi;
                g2[u][v] = 0ull;
            }
#pragma unroll
        for (int i = 0; i < 32; i++) { mine[kC1 * kC1 + (cg * 32 + i) * kC1 + ci] += (double)a[i]; a[i] = 0.f; }
        atomicAdd(mine + kC1 * kC1 + kC1 * kC2 + ci, (double)hs);
        hs = 0.f;
    };
    const long long ntiles = (P + kTP - 1) / kTP;
    int done = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        __syncthreads();
        load_x_tile<C0>(x0, sX, tile * kTP, P);
        for (int e = t; e < 4 * kC2; e += kThreads) {
            const long long m = tile * 4 + (e >> 7);
            sGs[e] = m < M ? __ldg(gs + m * kC2 + (e & 127)) : 0.f;
            sArg[e] = m < M ? __ldg(arg + m * kC2 + (e & 127)) : 0;
        }
        __syncthreads();
        h1_tile<C0>(sX, sH, w1, lp, ci, cg, tile * 4 + cg < M);
        __syncthreads();
#pragma unroll 4
        for (int pos = 0; pos < kTP; pos++) {
            const float4 hi = *reinterpret_cast<const float4 *>(sH + pos * kLdH + gi * 4);
            const float4 hj = *reinterpret_cast<const float4 *>(sH + pos * kLdH + gj * 4);
            const f2_t j01 = pack2(hj.x, hj.y), j23 = pack2(hj.z, hj.w);
            const f2_t ix = pack2(hi.x, hi.x), iy = pack2(hi.y, hi.y), iz = pack2(hi.z, hi.z), iw = pack2(hi.w, hi.w);
            ffma2(g2[0][0], ix, j01); ffma2(g2[0][1], ix, j23);
            ffma2(g2[1][0], iy, j01); ffma2(g2[1][1], iy, j23);
            ffma2(g2[2][0], iz, j01); ffma2(g2[2][1], iz, j23);
            ffma2(g2[3][0], iw, j01); ffma2(g2[3][1], iw, j23);
        }
#pragma unroll 8
        for (int r = 0; r < kK; r++) hs += sH[(cg * kK + r) * kLdH + ci];
#pragma unroll
        for (int mm = 0; mm < 4; mm++) {
#pragma unroll
            for (int cc = 0; cc < 32; cc++) {
                const float gv = sGs[mm * kC2 + cg * 32 + cc];      // warp-uniform
                if (gv != 0.f) a[cc] = __fmaf_rn(gv, sH[(mm * kK + sArg[mm * kC2 + cg * 32 + cc]) * kLdH + ci], a[cc]);
            }
        }
        if ((++done & 15) == 0) flush();
    }
    flush();
}

// g_h1 = (sparse rows) W2^T - h1 Q + cvec per 128-position tile; g_y1 = g_h1 (h1 > 0); s1_1 = sum g_y1,
// D = g_y1^T x0, t[p, f] = sum_c alpha[f][c] g_y1[p, c].  Thread tile 8 positions x 4 channels.
template <int C0>
__global__ void __launch_bounds__(kThreads, 2)
b_layer_kernel(long long P, int M, const float *__restrict__ x0, const float *__restrict__ flt,
               const float *__restrict__ Wt1, const float *__restrict__ W2T, const unsigned *__restrict__ lst,
               const float *__restrict__ lgs, const int *__restrict__ lcnt, float *__restrict__ tbuf, double *__restrict__ dbl) {
    constexpr int C_in = C0 - 3;
    extern __shared__ __align__(16) unsigned char tf_sm[];
    float *sH = reinterpret_cast<float *>(tf_sm);   // [128][68]
    float *sQ = sH + kTP * kLdH;                    // [64][64]
    float *sS = sQ + kC1 * kC1;                     // [128][64]
    float *sX = sS + kTP * kC1;                     // [128][C0] (room for 8)
    float *sLgs = sX + kTP * 8;                     // [4][128] list values (coefA g_y), sorted by sample
    unsigned *sLst = reinterpret_cast<unsigned *>(sLgs + 4 * kC2);   // [4][128] (sample << 8) | channel
    float *sAl = reinterpret_cast<float *>(sLst + 4 * kC2);  // [C_in][64]
    float *sCv = sAl + 5 * kC1;                     // [64]
    __shared__ int sCnt[4];
    const int t = threadIdx.x;
    for (int e = t; e < kC1 * kC1; e += kThreads) sQ[e] = __ldg(flt + fQ + e);
    for (int e = t; e < C_in * kC1; e += kThreads) sAl[e] = __ldg(flt + fAlpha + e);
    if (t < kC1) sCv[t] = __ldg(flt + fCvec + t);
    const int c1 = t & 63, q = t >> 6, pg = t >> 4, cg = t & 15;
    float w1[C0];
    L1Params lp;
    load_l1<C0>(Wt1, flt, c1, w1, lp);
    float s1a[4], Da[4][C0];
#pragma unroll
    for (int v = 0; v < 4; v++) {
        s1a[v] = 0.f;
#pragma unroll
        for (int i = 0; i < C0; i++) Da[v][i] = 0.f;
    }
    const long long ntiles = (P + kTP - 1) / kTP;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        __syncthreads();
        load_x_tile<C0>(x0, sX, tile * kTP, P);
        if (t < 4) sCnt[t] = tile * 4 + t < M ? __ldg(lcnt + tile * 4 + t) : 0;
        for (int e = t; e < 4 * kC2; e += kThreads) {       // entries past the count are never used
            const long long m = tile * 4 + (e >> 7);
            sLgs[e] = m < M ? __ldg(lgs + m * kC2 + (e & 127)) : 0.f;
            sLst[e] = m < M ? __ldg(lst + m * kC2 + (e & 127)) : 0u;
        }
        __syncthreads();
        h1_tile<C0>(sX, sH, w1, lp, c1, q, tile * 4 + q < M);
#pragma unroll 8
        for (int r = 0; r < kK; r++) sS[(q * kK + r) * kC1 + c1] = 0.f;
        {   // rows of the sparse term: the list is sorted by sample, so a row is accumulated in a register
            const int n = sCnt[q];
            int kcur = 0;
            float racc = 0.f;
            float *srow = sS + (q * kK) * kC1 + c1;
            for (int e = 0; e < n; e += 4) {
                const uint4 pk = *reinterpret_cast<const uint4 *>(sLst + q * kC2 + e);
                const float4 gv = *reinterpret_cast<const float4 *>(sLgs + q * kC2 + e);
                const float w0 = __ldg(W2T + (pk.x & 255u) * kC1 + c1), w1v = __ldg(W2T + (pk.y & 255u) * kC1 + c1);
                const float w2 = __ldg(W2T + (pk.z & 255u) * kC1 + c1), w3 = __ldg(W2T + (pk.w & 255u) * kC1 + c1);
                int k = (int)(pk.x >> 8);
                if (k != kcur) { srow[kcur * kC1] = racc; racc = 0.f; kcur = k; }
                racc = __fmaf_rn(gv.x, w0, racc);
                k = (int)(pk.y >> 8);
                if (k != kcur) { srow[kcur * kC1] = racc; racc = 0.f; kcur = k; }
                racc = __fmaf_rn(gv.y, w1v, racc);
                k = (int)(pk.z >> 8);
                if (k != kcur) { srow[kcur * kC1] = racc; racc = 0.f; kcur = k; }
                racc = __fmaf_rn(gv.z, w2, racc);
                k = (int)(pk.w >> 8);
                if (k != kcur) { srow[kcur * kC1] = racc; racc = 0.f; kcur = k; }
                racc = __fmaf_rn(gv.w, w3, racc);
            }
            srow[kcur * kC1] = racc;
        }
        __syncthreads();
        f2_t acc2[8][2];
#pragma unroll
        for (int r = 0; r < 8; r++) { acc2[r][0] = 0ull; acc2[r][1] = 0ull; }
        const float *aBase = sH + pg * kLdH;                // positions 16 r + pg: the two half-warps read adjacent rows
        const float *bBase = sQ + cg * 4;
#pragma unroll 1
        for (int k4 = 0; k4 < kC1 / 4; k4++) {
            float4 a[8];
#pragma unroll
            for (int r = 0; r < 8; r++) a[r] = *reinterpret_cast<const float4 *>(aBase + r * 16 * kLdH + k4 * 4);
#pragma unroll
            for (int kk = 0; kk < 4; kk++) {
                const float4 b = *reinterpret_cast<const float4 *>(bBase + (k4 * 4 + kk) * kC1);
                const f2_t b01 = pack2(b.x, b.y), b23 = pack2(b.z, b.w);
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    const float av = kk == 0 ? a[r].x : kk == 1 ? a[r].y : kk == 2 ? a[r].z : a[r].w;
                    const f2_t aa = pack2(av, av);
                    ffma2(acc2[r][0], aa, b01);
                    ffma2(acc2[r][1], aa, b23);
                }
            }
        }
        float acc[8][4];
#pragma unroll
        for (int r = 0; r < 8; r++) {
            unpack2(acc2[r][0], acc[r][0], acc[r][1]);
            unpack2(acc2[r][1], acc[r][2], acc[r][3]);
        }
        const float4 cv = *reinterpret_cast<const float4 *>(sCv + cg * 4);
        const long long p0 = tile * kTP;
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const int pos = r * 16 + pg;
            const float4 h4 = *reinterpret_cast<const float4 *>(sH + pos * kLdH + cg * 4);
            const float4 s4 = *reinterpret_cast<const float4 *>(sS + pos * kC1 + cg * 4);
            float gy[4];
            gy[0] = h4.x > 0.f ? (s4.x - acc[r][0]) + cv.x : 0.f;
            gy[1] = h4.y > 0.f ? (s4.y - acc[r][1]) + cv.y : 0.f;
            gy[2] = h4.z > 0.f ? (s4.z - acc[r][2]) + cv.z : 0.f;
            gy[3] = h4.w > 0.f ? (s4.w - acc[r][3]) + cv.w : 0.f;
            float x[C0];
#pragma unroll
            for (int i = 0; i < C0; i++) x[i] = sX[pos * C0 + i];
            float tt[C_in > 0 ? C_in : 1];
#pragma unroll
            for (int f = 0; f < C_in; f++) tt[f] = 0.f;
#pragma unroll
            for (int v = 0; v < 4; v++) {
                s1a[v] += gy[v];
#pragma unroll
                for (int i = 0; i < C0; i++) Da[v][i] = __fmaf_rn(gy[v], x[i], Da[v][i]);
#pragma unroll
                for (int f = 0; f < C_in; f++) tt[f] = __fmaf_rn(sAl[f * kC1 + cg * 4 + v], gy[v], tt[f]);
            }
            if (tbuf) {
#pragma unroll
                for (int f = 0; f < C_in; f++) {
#pragma unroll
                    for (int o = 1; o < 16; o <<= 1) tt[f] += __shfl_xor_sync(0xffffffffu, tt[f], o);
                    if (cg == 0 && p0 + pos < P) tbuf[(p0 + pos) * C_in + f] = tt[f];
                }
            }
        }
    }
#pragma unroll
    for (int v = 0; v < 4; v++) {
        s1a[v] += __shfl_xor_sync(0xffffffffu, s1a[v], 16);
#pragma unroll
        for (int i = 0; i < C0; i++) Da[v][i] += __shfl_xor_sync(0xffffffffu, Da[v][i], 16);
    }
    if ((t & 31) < 16) {
#pragma unroll
        for (int v = 0; v < 4; v++) {
            atomicAdd(dbl + dS11 + cg * 4 + v, (double)s1a[v]);
#pragma unroll
            for (int i = 0; i < C0; i++) atomicAdd(dbl + dD + (cg * 4 + v) * 8 + i, (double)Da[v][i]);
        }
    }
}

__global__ void __launch_bounds__(kThreads)
b_reduce_kernel(int n_cta, const double *__restrict__ part, double *__restrict__ red) {
    const int e = blockIdx.x * kThreads + threadIdx.x;
    if (e >= kStatEnt) return;
    double v = 0.0;
    for (int k = 0; k < n_cta; k++) v += part[(size_t)k * kStatEnt + e];
    red[e] = v;
}

// blocks 0..31: dW2^T; block 32: layer 1 (dW1^T, dgamma1, dbeta1) and the coefficients of the feature gradient
template <int C0>
__global__ void __launch_bounds__(kThreads)
b_finalize_kernel(long long P, const float *__restrict__ Wt1, const float *__restrict__ Wt2, float *__restrict__ flt,
                  const double *__restrict__ dbl, float *__restrict__ gW, float *__restrict__ gGB) {
    constexpr int C_in = C0 - 3;
    const int t = threadIdx.x;
    const double invP = 1.0 / (double)P;
    if (blockIdx.x < 32) {
        const int e = blockIdx.x * kThreads + t, ci = e >> 7, c = e & 127;
        const double *G = dbl + dRed, *AT = G + kC1 * kC1, *H = AT + kC1 * kC2;
        double gw = 0.0;
        for (int j = 0; j < kC1; j++) gw += G[ci * kC1 + j] * (double)__ldg(Wt2 + j * kC2 + c);
        const double val = AT[c * kC1 + ci] - (double)flt[fU + c] * H[ci] - (double)flt[fE + c] * (gw - (double)flt[fMean2 + c] * H[ci]);
        gW[C0 * kC1 + e] = (float)val;
        return;
    }
    __shared__ double s1s[kC1], s2s[kC1];
    if (t < kC1) {
        const double s1 = dbl[dS11 + t];
        double raw = 0.0;
#pragma unroll
        for (int i = 0; i < C0; i++) raw += (double)Wt1[i * kC1 + t] * dbl[dD + t * 8 + i];
        const double s2 = (double)flt[fInv1 + t] * (raw - (double)flt[fMean1 + t] * s1);
        s1s[t] = s1; s2s[t] = s2;
        gGB[t] = (float)s2;
        gGB[kC1 + t] = (float)s1;
    }
    __syncthreads();
    for (int e = t; e < C0 * kC1; e += kThreads) {
        const int i = e / kC1, ci = e - i * kC1;
        double gx = 0.0;
#pragma unroll
        for (int j = 0; j < C0; j++) gx += g0_sym(dbl, i, j) * (double)Wt1[j * kC1 + ci];
        const double X = dbl[dX + i], mean = (double)flt[fMean1 + ci], inv = (double)flt[fInv1 + ci];
        const double val = (double)flt[fSc1 + ci] * (dbl[dD + ci * 8 + i] - s1s[ci] * invP * X - s2s[ci] * invP * inv * (gx - mean * X));
        gW[e] = (float)val;
    }
    if (t < C_in * C0) {
        const int f = t / C0, i = t - f * C0;
        double qv = 0.0;
        for (int ci = 0; ci < kC1; ci++)
            qv += (double)flt[fSc1 + ci] * s2s[ci] * invP * (double)flt[fInv1 + ci] * (double)Wt1[i * kC1 + ci] * (double)Wt1[(3 + f) * kC1 + ci];
        flt[fQf + f * 8 + i] = (float)qv;
    } else if (t >= 64 && t < 64 + C_in) {
        const int f = t - 64;
        double cv = 0.0;
        for (int ci = 0; ci < kC1; ci++)
            cv += (double)flt[fSc1 + ci] * (double)Wt1[(3 + f) * kC1 + ci] * (s2s[ci] * invP * (double)flt[fInv1 + ci] * (double)flt[fMean1 + ci] - s1s[ci] * invP);
        flt[fCf + f] = (float)cv;
    }
}

// grad_features[src(p), f] += t[p, f] - x0[p] . q_f + c_f for every sample of a non-empty ball
template <int C0>
__global__ void __launch_bounds__(kThreads)
b_scatter_kernel(long long P, int B, const float *__restrict__ x0, const float *__restrict__ tbuf,
                 const float *__restrict__ flt, const int *__restrict__ pre_xyz, const int *__restrict__ pre_new,
                 int uniform, int n_u, int m_u, const int *__restrict__ idx, const unsigned char *__restrict__ empty_mask,
                 float *__restrict__ grad_features) {
    constexpr int C_in = C0 - 3;
    for (long long p = (long long)blockIdx.x * kThreads + threadIdx.x; p < P; p += (long long)gridDim.x * kThreads) {
        const int m = (int)(p >> 5);
        if ((__ldg(idx + (size_t)m * kK) < 0) || (empty_mask && empty_mask[m])) continue;
        const long long start = uniform ? (long long)(m / m_u) * n_u : __ldg(pre_xyz + find_batch(pre_new, B, m));
        const long long src = start + __ldg(idx + p);
        float x[C0];
#pragma unroll
        for (int i = 0; i < C0; i++) x[i] = __ldg(x0 + p * C0 + i);
#pragma unroll
        for (int f = 0; f < C_in; f++) {
            float v = __ldg(tbuf + p * C_in + f) + __ldg(flt + fCf + f);
#pragma unroll
            for (int i = 0; i < C0; i++) v = __fmaf_rn(-x[i], __ldg(flt + fQf + f * 8 + i), v);
            atomicAdd(grad_features + src * C_in + f, v);
        }
    }
}

constexpr size_t kSmemF1 = (size_t)(kC1 * kC2 + kTP * kLdH + kTP * 8) * 4 + 8 * 8 * 16 * 8;
constexpr size_t kSmemB1a = (size_t)(kTP * kLdH + kTP * 8 + 4 * kC2) * 4 + 4 * kC2 * 4;
constexpr size_t kSmemB1 = (size_t)(kTP * kLdH + kC1 * kC1 + kTP * kC1 + kTP * 8 + 4 * kC2) * 4 + 4 * kC2 * 4 + (5 * kC1 + kC1) * 4;

int grid_tiles(int M) {
    const long long tiles = ((long long)M + 3) / 4;
    return (int)(tiles < kMaxGrid ? tiles : kMaxGrid);
}

template <int C0>
int forward_t(int B, int M, const float *xyz, const float *new_xyz, const float *features, const int *pre_xyz,
              const int *pre_new, int uniform, int n_u, int m_u, const int *idx, const unsigned char *empty_mask,
              const float *weights, const float *gamma_beta, float eps, float *out, int ld_out, float *batch_stats,
              char *W, const Layout &l, cudaStream_t st) {
    const long long P = (long long)M * kK;
    float *x0 = reinterpret_cast<float *>(W + l.x0);
    double *dbl = reinterpret_cast<double *>(W + l.dbl);
    float *flt = reinterpret_cast<float *>(W + l.flt);
    float *zmax = reinterpret_cast<float *>(W + l.zmax);
    int *arg = reinterpret_cast<int *>(W + l.arg);
    MGAR_CUDA(cudaMemsetAsync(dbl, 0, (size_t)dFwdEnd * 8, st));
    int g0 = div_up(M, kThreads / 32);
    if (g0 > kNumSMs * 8) g0 = kNumSMs * 8;
    f_group_kernel<C0><<<g0, kThreads, 0, st>>>(B, M, xyz, new_xyz, features, pre_xyz, pre_new, uniform, n_u, m_u, idx,
                                                  empty_mask, x0, dbl);
    f_bn1_kernel<C0><<<1, kC1, 0, st>>>(P, eps, weights, gamma_beta, dbl, flt, batch_stats);
    const char *mode = getenv("MGAR_SA_TRAIN_KERNELS");
    const bool tc = mode && strcmp(mode, "tc") == 0;
    if (!tc) {
        MGAR_CUDA(cudaFuncSetAttribute(f_layer2_kernel<C0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemF1));
        f_layer2_kernel<C0><<<grid_tiles(M), kThreads, kSmemF1, st>>>(P, M, x0, flt, weights, weights + C0 * kC1, gamma_beta + 2 * kC1, zmax, arg, dbl);
        count_launch();
    } else {
        // layer 2 on the tensor cores: the inference kernel (sa_mlp_tc16.cu, tcgen05 3xFP16) with the statistics
        // epilogue.  Its layer 1 (W1 with BatchNorm 1's batch scale folded in, the shift, the bound of the pre-activation)
        // is read from the device (f_bn1_kernel wrote it behind the other coefficients): no copy back, no synchronisation.
        TcLayer1 L1 = {};
        TcArgs A;
        A.live_groups = nullptr; A.n_live = nullptr; A.live_tiles = nullptr; A.live_units = nullptr; A.m_shift = -1; A.ns = kTcNS;
        A.B = B; A.M = M; A.C_in = C0 - 3;
        A.xyz = xyz; A.xyz_stride = 3; A.new_xyz = new_xyz;
        A.features = features; A.feat_stride = C0 - 3;
        A.pre_xyz = pre_xyz; A.pre_new = pre_new; A.uniform = uniform; A.n_u = n_u; A.m_u = m_u;
        A.idx = idx; A.empty_mask = empty_mask;
        A.w2t = weights + C0 * kC1; A.sc2 = flt + fSgn2; A.sh2 = flt + fZero2;
        A.out = nullptr; A.ld_out = 0; A.gmax = nullptr; A.group_size = 1; A.group_zero = nullptr;
        A.packed4 = 0;
#ifdef MGAR_TC_TRACE
        A.trace = nullptr;
#endif
        A.tr_zext = zmax; A.tr_arg = arg; A.tr_sum = dbl + dSum2; A.tr_sq = dbl + dSq2;
        A.l1_dev = flt + fL1;
        const int n_tiles = (M + 3) / 4;
        if (int rc = sa_mlp_tc16_launch_train(L1, A, C0, n_tiles < kNumSMs ? n_tiles : kNumSMs, st)) return rc;
        int gf = div_up((long long)M * kC2, kThreads * 4);
        if (gf > kNumSMs * 8) gf = kNumSMs * 8;
        f_fix_ties_kernel<C0><<<gf, kThreads, 0, st>>>((long long)M * kC2, x0, flt, weights, weights + C0 * kC1, gamma_beta + 2 * kC1, zmax, arg);
        count_launch();
    }
    f_bn2_kernel<<<1, kC2, 0, st>>>(P, eps, gamma_beta, dbl, flt, batch_stats);
    f_pool_kernel<C0><<<div_up((long long)M * kC2, kThreads), kThreads, 0, st>>>((long long)M * kC2, flt, gamma_beta + 2 * kC1, zmax, out, ld_out,
                                                                                   tc ? x0 : nullptr, arg, weights, weights + C0 * kC1);
    count_launch(4);
    return check_launch("sa_mlp_fwd_train (fast)");
}

template <int C0>
int backward_t(int B, int M, const int *pre_xyz, const int *pre_new, int uniform, int n_u, int m_u, const int *idx,
               const unsigned char *empty_mask, const float *weights, const float *grad_out, int ld_grad, char *W,
               const Layout &l, float *gW, float *gGB, float *grad_features, cudaStream_t st) {
    const long long P = (long long)M * kK;
    const float *x0 = reinterpret_cast<const float *>(W + l.x0);
    double *dbl = reinterpret_cast<double *>(W + l.dbl);
    float *flt = reinterpret_cast<float *>(W + l.flt);
    const float *zsel = reinterpret_cast<const float *>(W + l.zmax);
    float *gs = reinterpret_cast<float *>(W + l.gs);
    const int *arg = reinterpret_cast<const int *>(W + l.arg);
    float *tbuf = grad_features ? reinterpret_cast<float *>(W + l.t) : nullptr;
    double *part = reinterpret_cast<double *>(W + l.part);
    float *W2T = reinterpret_cast<float *>(W + l.w2t);
    const float *Wt1 = weights, *Wt2 = weights + C0 * kC1;
    const int grid = grid_tiles(M);
    MGAR_CUDA(cudaMemsetAsync(dbl + dS1, 0, (size_t)(dRed - dS1) * 8, st));
    MGAR_CUDA(cudaMemsetAsync(part, 0, (size_t)grid * kStatEnt * 8, st));
    unsigned *lst = reinterpret_cast<unsigned *>(W + l.lst);
    float *lgs = reinterpret_cast<float *>(W + l.lgs);
    int *lcnt = reinterpret_cast<int *>(W + l.lcnt);
    int gg = div_up(M, kThreads / 32);
    if (gg > kNumSMs * 8) gg = kNumSMs * 8;
    b_gy_kernel<<<gg, kThreads, 0, st>>>(M, zsel, arg, flt, grad_out, ld_grad, gs, lst, lgs, lcnt, dbl);
    b_prep_kernel<<<17, kThreads, 0, st>>>(P, Wt2, dbl, flt, W2T, gGB);
    MGAR_CUDA(cudaFuncSetAttribute(b_h1stats_kernel<C0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemB1a));
    b_h1stats_kernel<C0><<<grid, kThreads, kSmemB1a, st>>>(P, M, x0, flt, Wt1, gs, arg, part);
    MGAR_CUDA(cudaFuncSetAttribute(b_layer_kernel<C0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemB1));
    b_layer_kernel<C0><<<grid, kThreads, kSmemB1, st>>>(P, M, x0, flt, Wt1, W2T, lst, lgs, lcnt, tbuf, dbl);
    b_reduce_kernel<<<div_up(kStatEnt, kThreads), kThreads, 0, st>>>(grid, part, dbl + dRed);
    b_finalize_kernel<C0><<<33, kThreads, 0, st>>>(P, Wt1, Wt2, flt, dbl, gW, gGB);
    count_launch(6);
    if (grad_features) {
        int gsz = div_up(P, kThreads);
        if (gsz > kNumSMs * 16) gsz = kNumSMs * 16;
        b_scatter_kernel<C0><<<gsz, kThreads, 0, st>>>(P, B, x0, tbuf, flt, pre_xyz, pre_new, uniform, n_u, m_u, idx, empty_mask,
                                                        grad_features);
        count_launch();
    }
    return check_launch("sa_mlp_bwd (fast)");
}

}  // namespace

bool applicable(int M, int nsample, int n_layers, const int *ch, int pool_avg) {
    if (pool_avg || n_layers != 2 || nsample != kK || !ch || M < 1) return false;
    if (ch[1] != kC1 || ch[2] != kC2 || (ch[0] != 4 && ch[0] != 5)) return false;
    const char *e = getenv("MGAR_SA_TRAIN_KERNELS");
    return !(e && strcmp(e, "generic") == 0);
}

size_t workspace_bytes(int M, int C0) { return make_layout(M, C0).total; }

int forward(int B, int M, int C_in, const float *xyz, const int *xyz_batch_cnt, const float *new_xyz,
            const int *new_xyz_batch_cnt, int uniform_n, int uniform_m, const float *features, const int *idx,
            const unsigned char *empty_mask, const float *weights, const float *gamma_beta, float eps, float *out,
            int ld_out, float *batch_stats, void *workspace, size_t ws_bytes, cudaStream_t st) {
    const int C0 = 3 + C_in;
    const Layout l = make_layout(M, C0);
    if (ws_bytes < l.total) { set_error("sa_mlp_train: workspace %zu < %zu bytes", ws_bytes, l.total); return MGAR_ENOMEM; }
    const bool uniform = uniform_n > 0;
    PrefixWs ws;
    if (!uniform) {   // ragged batches: the prefix arrays sit behind the training workspace
        const size_t off = (l.total + 15) & ~(size_t)15;
        if (int rc = ws.bind(ws_bytes > off ? static_cast<char *>(workspace) + off : nullptr, ws_bytes > off ? ws_bytes - off : 0, B, "sa_mlp_fwd_train")) return rc;
        if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    }
    char *W = static_cast<char *>(workspace);
    if (C0 == 4)
        return forward_t<4>(B, M, xyz, new_xyz, features, ws.b, ws.a, uniform ? 1 : 0, uniform_n, uniform_m, idx, empty_mask,
                            weights, gamma_beta, eps, out, ld_out, batch_stats, W, l, st);
    return forward_t<5>(B, M, xyz, new_xyz, features, ws.b, ws.a, uniform ? 1 : 0, uniform_n, uniform_m, idx, empty_mask,
                        weights, gamma_beta, eps, out, ld_out, batch_stats, W, l, st);
}

int backward(int B, int M, int C_in, const int *xyz_batch_cnt, const int *new_xyz_batch_cnt, int uniform_n,
             int uniform_m, const int *idx, const unsigned char *empty_mask, const float *weights,
             const float *gamma_beta, const float *grad_out, int ld_grad, void *workspace, size_t ws_bytes,
             float *grad_weights, float *grad_gamma_beta, float *grad_features, cudaStream_t st) {
    (void)gamma_beta;
    const int C0 = 3 + C_in;
    const Layout l = make_layout(M, C0);
    if (ws_bytes < l.total) { set_error("sa_mlp_bwd: workspace %zu < %zu bytes", ws_bytes, l.total); return MGAR_ENOMEM; }
    const bool uniform = uniform_n > 0;
    PrefixWs ws;
    if (!uniform && grad_features) {
        const size_t off = (l.total + 15) & ~(size_t)15;
        if (int rc = ws.bind(ws_bytes > off ? static_cast<char *>(workspace) + off : nullptr, ws_bytes > off ? ws_bytes - off : 0, B, "sa_mlp_bwd")) return rc;
        if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    }
    char *W = static_cast<char *>(workspace);
    if (C0 == 4)
        return backward_t<4>(B, M, ws.b, ws.a, uniform ? 1 : 0, uniform_n, uniform_m, idx, empty_mask, weights, grad_out, ld_grad,
                             W, l, grad_weights, grad_gamma_beta, grad_features, st);
    return backward_t<5>(B, M, ws.b, ws.a, uniform ? 1 : 0, uniform_n, uniform_m, idx, empty_mask, weights, grad_out, ld_grad,
                         W, l, grad_weights, grad_gamma_beta, grad_features, st);
}

}  // namespace trfast
}  // namespace mgar
