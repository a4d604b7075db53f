// Set-abstraction shared MLP, fused with grouping and the max-pool (fp32 CUDA-core path).
//
// Reference: StackSAModuleMSG.forward (pointnet2_stack/pointnet2_modules.py:78-112) after
// QueryAndGroup (pointnet2_utils.py:123-159): two gathers, a subtract, a mask, a cat, then
// [Conv2d 1x1, BatchNorm2d, ReLU] x L through cuDNN over materialised (C, M, K) tensors and a
// max_pool2d - about 2 GB of HBM traffic per JRDB frame for what is 10 MB of output.
//
// Here one kernel reads idx, gathers (xyz - centre | features) straight into shared memory, runs
// every layer with activations resident in shared memory (BatchNorm folded to scale/shift, eval
// mode), and reduces max over the K samples with warp REDUX before anything is written: the only
// HBM write is the (M, C_last) result.  A warp owns a centroid, a lane owns a sample.
// This is the exact-fp32 path (FFMA); the tcgen05 path for the 64->128 layer lives in
// sa_mlp_tc.cu and is validated against this one.
#include "mgar_common.cuh"
#include "stack_prefix.cuh"

namespace mgar {

constexpr int kMlpWarps = 8;
constexpr int kMaxLayers = 6;

struct MlpSpec {
    int n_layers;
    int ch[kMaxLayers + 1];     // ch[0] = 3 + C_in
    int w_off[kMaxLayers];      // float offsets into the packed parameter block
    int s_off[kMaxLayers];      // offsets of scale (then shift) per layer
    int hid;                    // widest hidden activation (+1 padding)
    int in_pad;                 // ch[0] padded to odd
};

__device__ __forceinline__ int find_batch_mlp(const int *__restrict__ pre, int B, int v) {
    int lo = 0, hi = B;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(pre + mid) <= v) lo = mid; else hi = mid;
    }
    return lo;
}

// wt: per layer W^T laid out [c_in][c_out] (so a lane-uniform c reads a contiguous row: LDS.128
// broadcasts), scale/shift: folded BN.  All staged in shared memory once per CTA.
__global__ void __launch_bounds__(kMlpWarps * 32)
sa_mlp_fp32_kernel(MlpSpec sp, int B, int M, int K, int C_in, const float *__restrict__ xyz,
                   const float *__restrict__ new_xyz, const float *__restrict__ features,
                   const int *__restrict__ pre_xyz, const int *__restrict__ pre_new, int uniform, int n_u, int m_u,
                   const int *__restrict__ idx, const unsigned char *__restrict__ empty_mask,
                   const float *__restrict__ params, int n_params, float *__restrict__ out, int ld_out,
                   int pool_avg) {
    extern __shared__ float smem[];
    float *s_par = smem;                                  // n_params floats
    const int per_warp = 32 * sp.in_pad + 2 * 32 * sp.hid;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *s_in = smem + ((n_params + 3) & ~3) + warp * per_warp;   // [32][in_pad]
    float *s_h0 = s_in + 32 * sp.in_pad;                              // [32][hid]
    float *s_h1 = s_h0 + 32 * sp.hid;

    for (int i = threadIdx.x; i < n_params; i += kMlpWarps * 32) s_par[i] = __ldg(params + i);
    __syncthreads();

    for (int m = blockIdx.x * kMlpWarps + warp; m < M; m += gridDim.x * kMlpWarps) {
        long long start;
        if (uniform) start = (long long)(m / m_u) * n_u;
        else start = __ldg(pre_xyz + find_batch_mlp(pre_new, B, m));
        const bool empty = (__ldg(idx + (size_t)m * K) < 0) || (empty_mask && empty_mask[m]);
        const float qx = __ldg(new_xyz + (size_t)m * 3 + 0), qy = __ldg(new_xyz + (size_t)m * 3 + 1),
                    qz = __ldg(new_xyz + (size_t)m * 3 + 2);
        // K may exceed 32: samples go through in passes of 32, pooled through out[]
        for (int s0 = 0; s0 < K; s0 += 32) {
            const int s = s0 + lane;
            const bool live = s < K;
            // ---- gather prologue: (xyz - centre | features), zeros for empty balls ---------------
            {
                float *row = s_in + lane * sp.in_pad;
                if (live && !empty) {
                    const long long p = start + __ldg(idx + (size_t)m * K + s);
                    row[0] = __fsub_rn(__ldg(xyz + p * 3 + 0), qx);
                    row[1] = __fsub_rn(__ldg(xyz + p * 3 + 1), qy);
                    row[2] = __fsub_rn(__ldg(xyz + p * 3 + 2), qz);
                    for (int c = 0; c < C_in; c++) row[3 + c] = __ldg(features + p * C_in + c);
                } else {
                    for (int c = 0; c < 3 + C_in; c++) row[c] = 0.f;
                }
            }
            __syncwarp();
            const float *a_in = s_in;
            int a_ld = sp.in_pad;
            for (int l = 0; l < sp.n_layers; l++) {
                const int ci = sp.ch[l], co = sp.ch[l + 1];
                const float *wt = s_par + sp.w_off[l];
                const float *sc = s_par + sp.s_off[l], *sh = sc + co;
                const bool last = (l == sp.n_layers - 1);
                float *a_out = (l & 1) ? s_h1 : s_h0;
                const float *arow = a_in + lane * a_ld;
                for (int ob = 0; ob < co; ob += 8) {
                    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    if (ob + 8 <= co && (co & 3) == 0) {
                        for (int c = 0; c < ci; c++) {
                            const float a = arow[c];
                            const float4 w0 = *reinterpret_cast<const float4 *>(wt + c * co + ob);
                            const float4 w1 = *reinterpret_cast<const float4 *>(wt + c * co + ob + 4);
                            acc[0] = __fmaf_rn(a, w0.x, acc[0]); acc[1] = __fmaf_rn(a, w0.y, acc[1]);
                            acc[2] = __fmaf_rn(a, w0.z, acc[2]); acc[3] = __fmaf_rn(a, w0.w, acc[3]);
                            acc[4] = __fmaf_rn(a, w1.x, acc[4]); acc[5] = __fmaf_rn(a, w1.y, acc[5]);
                            acc[6] = __fmaf_rn(a, w1.z, acc[6]); acc[7] = __fmaf_rn(a, w1.w, acc[7]);
                        }
                    } else {
                        for (int c = 0; c < ci; c++) {
                            const float a = arow[c];
#pragma unroll
                            for (int j = 0; j < 8; j++)
                                if (ob + j < co) acc[j] = __fmaf_rn(a, wt[c * co + ob + j], acc[j]);
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        if (ob + j >= co) break;
                        float y = fmaxf(__fmaf_rn(acc[j], sc[ob + j], sh[ob + j]), 0.f);
                        if (!last) {
                            a_out[lane * sp.hid + ob + j] = y;
                        } else {
                            // pool over the samples of this pass; y >= 0 so uint order == float order
                            float r;
                            if (pool_avg) {
                                y = live ? y : 0.f;
#pragma unroll
                                for (int d = 16; d >= 1; d >>= 1) y += __shfl_xor_sync(0xffffffffu, y, d);
                                r = y;
                            } else {
                                r = __uint_as_float(__reduce_max_sync(0xffffffffu, live ? __float_as_uint(y) : 0u));
                            }
                            if (lane == 0) {
                                float *o = out + (size_t)m * ld_out + ob + j;
                                if (s0 == 0) *o = r;
                                else *o = pool_avg ? *o + r : fmaxf(*o, r);
                                if (pool_avg && s0 + 32 >= K) *o = *o / (float)K;
                            }
                        }
                    }
                }
                __syncwarp();
                a_in = a_out;
                a_ld = sp.hid;
            }
        }
    }
}

}  // namespace mgar

using namespace mgar;

// Parameter block layout (device, floats): for each layer l: W^T [ch[l]][ch[l+1]], then
// scale[ch[l+1]], shift[ch[l+1]] (BatchNorm folded; the Python host layer packs it).
extern "C" int mgar_sa_mlp_fwd_fp32(int B, int M, int nsample, int C_in, const float *xyz, const int *xyz_batch_cnt,
                                    const float *new_xyz, const int *new_xyz_batch_cnt, int uniform_n,
                                    int uniform_m, const float *features, const int *idx,
                                    const unsigned char *empty_mask, int n_layers, const int *channels_host,
                                    const float *params, float *out, int ld_out, int pool_avg, void *workspace,
                                    size_t workspace_bytes, void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0 && nsample >= 1 && C_in >= 0, "sa_mlp: bad size");
    MGAR_REQUIRE(n_layers >= 1 && n_layers <= kMaxLayers, "sa_mlp: 1..%d layers supported", kMaxLayers);
    if (M == 0) return MGAR_OK;
    MGAR_REQUIRE(xyz && new_xyz && idx && params && out && channels_host, "sa_mlp: null pointer");
    MGAR_REQUIRE(features || C_in == 0, "sa_mlp: features missing");
    const bool uniform = uniform_n > 0;
    MGAR_REQUIRE(uniform || (xyz_batch_cnt && new_xyz_batch_cnt), "sa_mlp: batch counts missing");
    MGAR_REQUIRE(channels_host[0] == 3 + C_in, "sa_mlp: channels[0]=%d but 3+C_in=%d", channels_host[0], 3 + C_in);
    MlpSpec sp;
    sp.n_layers = n_layers;
    int off = 0, hid = 1;
    for (int l = 0; l <= n_layers; l++) {
        MGAR_REQUIRE(channels_host[l] >= 1, "sa_mlp: bad channel count");
        sp.ch[l] = channels_host[l];
    }
    for (int l = 0; l < n_layers; l++) {
        sp.w_off[l] = off;
        off += sp.ch[l] * sp.ch[l + 1];
        sp.s_off[l] = off;
        off += 2 * sp.ch[l + 1];
        if (l < n_layers - 1 && sp.ch[l + 1] > hid) hid = sp.ch[l + 1];
    }
    sp.hid = hid | 1;
    sp.in_pad = sp.ch[0] | 1;
    const int per_warp = 32 * sp.in_pad + 2 * 32 * sp.hid;
    const size_t smem = ((size_t)((off + 3) & ~3) + (size_t)kMlpWarps * per_warp) * sizeof(float);
    MGAR_REQUIRE(smem <= 220 * 1024, "sa_mlp: layer spec needs %zu bytes of shared memory", smem);
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (!uniform) {
        if (int rc = ws.bind(workspace, workspace_bytes, B, "sa_mlp_fwd_fp32")) return rc;
        if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    }
    MGAR_CUDA(cudaFuncSetAttribute(sa_mlp_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = div_up(M, kMlpWarps);
    const int cap = kNumSMs * 2;  // persistent-style: at most 2 CTAs per SM, parameters staged once each
    if (grid > cap) grid = cap;
    sa_mlp_fp32_kernel<<<grid, kMlpWarps * 32, smem, st>>>(sp, B, M, nsample, C_in, xyz, new_xyz, features, ws.b, ws.a,
                                                           uniform ? 1 : 0, uniform_n, uniform_m, idx, empty_mask,
                                                           params, off, out, ld_out, pool_avg);
    count_launch();
    return check_launch("sa_mlp_fp32");
}
