// NeighborVoxelSAModuleMSG, one scale, eval mode, as two kernels (SURVEY 8 f1 - the pooling module mil3.yaml runs).
//
// Reference: pcdet/ops/pointnet2/pointnet2_stack/voxel_pool_modules.py:70-130.  Per scale the module computes
//   f_in  = BN1d(Conv1d(features))                      on all N voxel centres   (mlps_in,  C_in -> C1, no ReLU)
//   group = f_in[idx] , offs = xyz[idx] - new_xyz       (M, ns) neighbours from the voxel query
//   h     = ReLU(group + BN2d(Conv2d(offs)))            (mlps_pos, 3 -> C1)
//   p     = max / mean over the ns samples
//   out   = ReLU(BN1d(Conv1d(p)))                       (mlps_out, C1 -> C2)
// through cuDNN / ATen with four materialised (M, C, ns) tensors.  The widths are small (mil3.yaml: 32 -> 32 -> 32,
// ns = 16): a gather-bound op whose contractions are 32-wide - they run here in exact fp32 FFMA on the CUDA cores,
// fused with the gather (a 128 x 128 tensor-core tile would be 94 % padding), BatchNorm folded to scale / shift.
//   kernel 1 (vsa_in_kernel):   warp per voxel centre, lanes = output channels, the feature row broadcast by shuffles.
//   kernel 2 (vsa_pool_kernel): warp per query point: lanes = channels; per sample one 128-byte gather of f_in, the
//                               3 -> C1 position term, ReLU, running max / sum; then the C1 -> C2 output layer with the
//                               pooled vector in shared memory.  Nothing of size (M, C, ns) is ever written.
#include <mutex>

#include "mgar_common.cuh"

namespace mgar {
namespace {

constexpr int kVsaWarps = 8;
constexpr int kVsaMaxC = 128;                 // C_in, C1, C2 <= 128 (<= 4 channels per lane)

// params layout (floats): Win^T [C_in][C1] | sc_in [C1] | sh_in [C1] | Wpos^T [3][C1] | sc_pos | sh_pos | Wout^T [C1][C2] | sc_out [C2] | sh_out [C2]
struct VsaOff { int win, sci, shi, wpos, scp, shp, wout, sco, sho, total; };
__host__ __device__ inline VsaOff vsa_off(int C_in, int C1, int C2) {
    VsaOff o;
    o.win = 0; o.sci = C_in * C1; o.shi = o.sci + C1; o.wpos = o.shi + C1; o.scp = o.wpos + 3 * C1; o.shp = o.scp + C1;
    o.wout = o.shp + C1; o.sco = o.wout + C1 * C2; o.sho = o.sco + C2; o.total = o.sho + C2;
    return o;
}

__global__ void __launch_bounds__(kVsaWarps * 32)
vsa_in_kernel(int N, int C_in, int C1, const float *__restrict__ feat, const float *__restrict__ params, float *__restrict__ f_in) {
    extern __shared__ float vsa_sm[];                        // Win^T [C_in][C1] | sc | sh
    const VsaOff o = vsa_off(C_in, C1, 0);
    for (int i = threadIdx.x; i < o.wpos; i += blockDim.x) vsa_sm[i] = __ldg(params + i);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int n = blockIdx.x * kVsaWarps + warp; n < N; n += gridDim.x * kVsaWarps) {
        float f[kVsaMaxC / 32];
#pragma unroll
        for (int q = 0; q < kVsaMaxC / 32; q++) f[q] = (q * 32 + lane < C_in) ? __ldg(feat + (size_t)n * C_in + q * 32 + lane) : 0.f;
        float acc[kVsaMaxC / 32] = {0.f, 0.f, 0.f, 0.f};
        for (int c = 0; c < C_in; c++) {
            float fc = 0.f;
#pragma unroll
            for (int q = 0; q < kVsaMaxC / 32; q++)
                if ((c >> 5) == q) fc = __shfl_sync(0xffffffffu, f[q], c & 31);
#pragma unroll
            for (int q = 0; q < kVsaMaxC / 32; q++)
                if (q * 32 + lane < C1) acc[q] = __fmaf_rn(vsa_sm[c * C1 + q * 32 + lane], fc, acc[q]);
        }
#pragma unroll
        for (int q = 0; q < kVsaMaxC / 32; q++) {
            const int c1 = q * 32 + lane;
            if (c1 < C1) f_in[(size_t)n * C1 + c1] = __fmaf_rn(acc[q], vsa_sm[o.sci + c1], vsa_sm[o.shi + c1]);
        }
    }
}

__global__ void __launch_bounds__(kVsaWarps * 32)
vsa_pool_kernel(int M, int C_in, int C1, int C2, int ns, int pool_avg, const float *__restrict__ f_in,
                const float *__restrict__ xyz, const float *__restrict__ new_xyz, const int *__restrict__ idx,
                const unsigned char *__restrict__ empty, const float *__restrict__ params, float *__restrict__ out, int ld_out) {
    extern __shared__ float vsa_sm[];                        // params from wpos on | per-warp pooled vectors [warps][C1]
    const VsaOff o = vsa_off(C_in, C1, C2);
    const int np = o.total - o.wpos;
    for (int i = threadIdx.x; i < np; i += blockDim.x) vsa_sm[i] = __ldg(params + o.wpos + i);
    __syncthreads();
    const float *wpos = vsa_sm, *scp = vsa_sm + (o.scp - o.wpos), *shp = vsa_sm + (o.shp - o.wpos);
    const float *wout = vsa_sm + (o.wout - o.wpos), *sco = vsa_sm + (o.sco - o.wpos), *sho = vsa_sm + (o.sho - o.wpos);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *pooled = vsa_sm + np + warp * C1;
    for (int m = blockIdx.x * kVsaWarps + warp; m < M; m += gridDim.x * kVsaWarps) {
        const bool gone = empty && empty[m];
        const float qx = __ldg(new_xyz + (size_t)m * 3), qy = __ldg(new_xyz + (size_t)m * 3 + 1), qz = __ldg(new_xyz + (size_t)m * 3 + 2);
        float best[kVsaMaxC / 32];
#pragma unroll
        for (int q = 0; q < kVsaMaxC / 32; q++) best[q] = pool_avg ? 0.f : -INFINITY;
        for (int s0 = 0; s0 < ns; s0 += 32) {
            const int mine = (s0 + lane < ns) ? __ldg(idx + (size_t)m * ns + s0 + lane) : 0;
            const int cnt = min(32, ns - s0);
            for (int s = 0; s < cnt; s++) {
                const int j = __shfl_sync(0xffffffffu, mine, s);
                float ox = 0.f, oy = 0.f, oz = 0.f;
                if (!gone) {
                    ox = __fsub_rn(__ldg(xyz + (size_t)j * 3), qx);
                    oy = __fsub_rn(__ldg(xyz + (size_t)j * 3 + 1), qy);
                    oz = __fsub_rn(__ldg(xyz + (size_t)j * 3 + 2), qz);
                }
#pragma unroll
                for (int q = 0; q < kVsaMaxC / 32; q++) {
                    const int c1 = q * 32 + lane;
                    if (c1 < C1) {
                        float p = __fmul_rn(wpos[c1], ox);
                        p = __fmaf_rn(wpos[C1 + c1], oy, p);
                        p = __fmaf_rn(wpos[2 * C1 + c1], oz, p);
                        p = __fmaf_rn(p, scp[c1], shp[c1]);
                        const float g = gone ? 0.f : __ldg(f_in + (size_t)j * C1 + c1);
                        const float h = fmaxf(__fadd_rn(g, p), 0.f);
                        best[q] = pool_avg ? __fadd_rn(best[q], h) : fmaxf(best[q], h);
                    }
                }
            }
        }
        __syncwarp();
#pragma unroll
        for (int q = 0; q < kVsaMaxC / 32; q++) {
            const int c1 = q * 32 + lane;
            if (c1 < C1) pooled[c1] = pool_avg ? __fdiv_rn(best[q], (float)ns) : best[q];
        }
        __syncwarp();
        for (int c2 = lane; c2 < C2; c2 += 32) {
            float acc = 0.f;
            for (int c1 = 0; c1 < C1; c1++) acc = __fmaf_rn(wout[c1 * C2 + c2], pooled[c1], acc);
            out[(size_t)m * ld_out + c2] = fmaxf(__fmaf_rn(acc, sco[c2], sho[c2]), 0.f);
        }
    }
}

}  // namespace
}  // namespace mgar

using namespace mgar;

extern "C" int mgar_voxel_sa_supported(int C_in, int C1, int C2, int nsample) {
    return C_in >= 1 && C_in <= kVsaMaxC && C1 >= 1 && C1 <= kVsaMaxC && C2 >= 1 && C2 <= kVsaMaxC && nsample >= 1;
}
extern "C" size_t mgar_voxel_sa_param_floats(int C_in, int C1, int C2) { return (size_t)vsa_off(C_in, C1, C2).total; }
extern "C" size_t mgar_voxel_sa_workspace_bytes(int N, int C1) { return N > 0 && C1 > 0 ? (size_t)N * C1 * sizeof(float) : 0; }

extern "C" int mgar_voxel_sa_fwd(int N, int M, int C_in, int C1, int C2, int nsample, int pool_avg, const float *features,
                                 const float *xyz, const float *new_xyz, const int *idx, const unsigned char *empty_mask,
                                 const float *params, float *out, int ld_out, void *workspace, size_t workspace_bytes,
                                 void *stream) {
    MGAR_REQUIRE(N >= 0 && M >= 0, "voxel_sa: negative size");
    MGAR_REQUIRE(mgar_voxel_sa_supported(C_in, C1, C2, nsample), "voxel_sa: channels must be in [1, %d] (C_in=%d C1=%d C2=%d)",
                 kVsaMaxC, C_in, C1, C2);
    if (M == 0) return MGAR_OK;
    MGAR_REQUIRE(features && xyz && new_xyz && idx && params && out, "voxel_sa: null pointer");
    MGAR_REQUIRE(ld_out >= C2, "voxel_sa: ld_out < C2");
    if (!workspace || workspace_bytes < mgar_voxel_sa_workspace_bytes(N, C1)) {
        set_error("voxel_sa: needs a workspace of mgar_voxel_sa_workspace_bytes(N, C1) = %zu bytes", mgar_voxel_sa_workspace_bytes(N, C1));
        return MGAR_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    float *f_in = static_cast<float *>(workspace);
    const VsaOff o = vsa_off(C_in, C1, C2);
    if (N > 0) {
        const size_t sm = (size_t)o.wpos * sizeof(float);
        if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(vsa_in_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        int grid = div_up(N, kVsaWarps);
        if (grid > kNumSMs * 8) grid = kNumSMs * 8;
        vsa_in_kernel<<<grid, kVsaWarps * 32, sm, st>>>(N, C_in, C1, features, params, f_in);
        count_launch();
        if (int rc = check_launch("voxel_sa(in)")) return rc;
    }
    const size_t sm = (size_t)(o.total - o.wpos + kVsaWarps * C1) * sizeof(float);
    if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(vsa_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    int grid = div_up(M, kVsaWarps);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    vsa_pool_kernel<<<grid, kVsaWarps * 32, sm, st>>>(M, C_in, C1, C2, nsample, pool_avg, f_in, xyz, new_xyz, idx, empty_mask,
                                                      params, out, ld_out);
    count_launch();
    return check_launch("voxel_sa(pool)");
}

// ---- the same two kernels for callers that cannot allocate device memory ---------------------------------------------
// No counterpart in the reference: added for code under CUDA graph capture, tight
// serving loops and harnesses that forbid cudaMalloc.  The (N, C1) output of the input layer - the only scratch of this
// op - lives in a file-scope device array of MGAR_VOXEL_SA_STATIC_MAX_FLOATS floats (8 MiB) that every GPU loading the
// library holds until exit.  Host code cannot use the array's name as a pointer (it is the host-side shadow of the
// symbol): the device address is taken with cudaGetSymbolAddress, once per device, with that device current.
// Every call on a device shares the array: vsa_in_kernel overwrites rows [0, N) before vsa_pool_kernel reads them and
// nothing needs to start at zero (rows >= N keep the previous call's values and are never read); two calls that overlap
// on the GPU (different streams, no dependency) race on it.
namespace mgar {
namespace {
constexpr size_t kVsaStaticFloats = MGAR_VOXEL_SA_STATIC_MAX_FLOATS;
static __device__ __align__(256) float g_vsa_static[kVsaStaticFloats];
constexpr int kVsaMaxDevices = 64;
float *g_vsa_static_ptr[kVsaMaxDevices];                      // device address of g_vsa_static, per device
std::mutex g_vsa_static_mu;
}  // namespace
}  // namespace mgar

extern "C" int mgar_voxel_sa_fwd_static(int N, int M, int C_in, int C1, int C2, int nsample, int pool_avg, const float *features,
                                        const float *xyz, const float *new_xyz, const int *idx,
                                        const unsigned char *empty_mask, const float *params, float *out, int ld_out,
                                        void *stream) {
    MGAR_REQUIRE(N >= 0 && M >= 0, "voxel_sa_static: negative size");
    MGAR_REQUIRE(mgar_voxel_sa_supported(C_in, C1, C2, nsample), "voxel_sa_static: channels must be in [1, %d] (C_in=%d C1=%d C2=%d)",
                 kVsaMaxC, C_in, C1, C2);
    if ((size_t)N * (size_t)C1 > kVsaStaticFloats) {
        set_error("voxel_sa_static: N * C1 = %zu floats exceeds the built-in scratch of %zu floats (%zu MiB); use "
                  "mgar_voxel_sa_fwd with a caller-owned workspace", (size_t)N * (size_t)C1, kVsaStaticFloats,
                  kVsaStaticFloats * sizeof(float) >> 20);
        return MGAR_ENOMEM;
    }
    if (M == 0) return MGAR_OK;
    MGAR_REQUIRE(features && xyz && new_xyz && idx && params && out, "voxel_sa_static: null pointer");
    MGAR_REQUIRE(ld_out >= C2, "voxel_sa_static: ld_out < C2");
    int dev = -1;
    MGAR_CUDA(cudaGetDevice(&dev));
    MGAR_REQUIRE(dev >= 0 && dev < kVsaMaxDevices, "voxel_sa_static: device ordinal %d out of range", dev);
    float *f_in = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_vsa_static_mu);
        if (!g_vsa_static_ptr[dev]) {                         // set-up: once per device, that device current
            void *p = nullptr;
            MGAR_CUDA(cudaGetSymbolAddress(&p, g_vsa_static));
            MGAR_REQUIRE(p != nullptr, "voxel_sa_static: cudaGetSymbolAddress returned a null address");
            g_vsa_static_ptr[dev] = static_cast<float *>(p);
        }
        f_in = g_vsa_static_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const VsaOff o = vsa_off(C_in, C1, C2);
    if (N > 0) {
        const size_t sm = (size_t)o.wpos * sizeof(float);
        if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(vsa_in_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        int grid = div_up(N, kVsaWarps);
        if (grid > kNumSMs * 8) grid = kNumSMs * 8;
        vsa_in_kernel<<<grid, kVsaWarps * 32, sm, st>>>(N, C_in, C1, features, params, f_in);
        count_launch();
        if (int rc = check_launch("voxel_sa_static(in)")) return rc;
    }
    const size_t sm = (size_t)(o.total - o.wpos + kVsaWarps * C1) * sizeof(float);
    if (sm > 48 * 1024) MGAR_CUDA(cudaFuncSetAttribute(vsa_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    int grid = div_up(M, kVsaWarps);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    vsa_pool_kernel<<<grid, kVsaWarps * 32, sm, st>>>(M, C_in, C1, C2, nsample, pool_avg, f_in, xyz, new_xyz, idx, empty_mask,
                                                      params, out, ld_out);
    count_launch();
    return check_launch("voxel_sa_static(pool)");
}
