// pointnet2_batch surface: the channel-major (B,C,N) gather / scatter ops and the three-nearest
// interpolation of PointNet++ feature propagation.  Ball query and FPS of this module are the stack kernels with a
// uniform layout (stack_ops.cu, fps.cu); gather_points is group_points with one sample per row.
//
// Reference: pointnet2_batch/src/group_points_gpu.cu:13-92, sampling_gpu.cu:14-85, interpolate_gpu.cu:14-168 - one
// thread per OUTPUT ELEMENT with the channel in blockIdx.y: every index (and every interpolation weight) is
// re-read C times, and three_nn streams the known points from global memory once per thread.
// Here a thread owns one (row, sample) position, reads its index / weights once and walks a chunk of channels
// (writes stay coalesced: consecutive threads are consecutive positions of one channel plane); three_nn stages the
// known points in shared memory tiles.
#include "mgar_common.cuh"

namespace mgar {

constexpr int kBtThreads = 256;
constexpr int kBtChan = 8;        // channels per CTA row: grid.y = ceil(C / kBtChan)
constexpr int kBtChanGrad = 2;    // interpolation scatter (3 atomics per channel): more threads in flight pay (measured)
constexpr int kNnTile = 1024;     // known points per shared-memory tile (16 KB as float4)

// out (B,C,P) = points (B,C,N) gathered at idx (B,P) (P = npoints * nsample, flattened)
template <bool kGrad>
__global__ void __launch_bounds__(kBtThreads)
batch_gather_kernel(int C, int N, int P, const float *__restrict__ src, const int *__restrict__ idx, float *__restrict__ dst) {
    const int b = blockIdx.z, p = blockIdx.x * kBtThreads + threadIdx.x;
    if (p >= P) return;
    const int i = __ldg(idx + (size_t)b * P + p);
    const int c0 = blockIdx.y * kBtChan, c1 = min(C, c0 + kBtChan);
    for (int c = c0; c < c1; ++c) {
        const size_t plane = (size_t)b * C + c;
        if (kGrad) atomicAdd(dst + plane * N + i, __ldg(src + plane * P + p));   // src = grad_out (B,C,P), dst = grad_points (B,C,N)
        else dst[plane * P + p] = __ldg(src + plane * N + i);
    }
}

// three nearest known points of every unknown point: squared distances ascending + indices.  Strict '<' against
// the running bests while walking k upward: ties keep the lowest index, as the reference's serial scan does.
constexpr int kNnThreads = 128;
__global__ void __launch_bounds__(kNnThreads)
three_nn_kernel(int n, int m, const float *__restrict__ unknown, const float *__restrict__ known,
                float *__restrict__ dist2, int *__restrict__ idx) {
    __shared__ float4 sk[kNnTile];                               // one 128-bit broadcast load per candidate
    const int b = blockIdx.y, p = blockIdx.x * kNnThreads + threadIdx.x;
    const bool live = p < n;
    const size_t row = (size_t)b * n + (live ? p : 0);
    const float ux = unknown[row * 3], uy = unknown[row * 3 + 1], uz = unknown[row * 3 + 2];
    known += (size_t)b * m * 3;
    float b1 = INFINITY, b2 = INFINITY, b3 = INFINITY;          // the reference's 1e40 doubles: above every float
    int i1 = 0, i2 = 0, i3 = 0;
    for (int k0 = 0; k0 < m; k0 += kNnTile) {
        const int kt = min(kNnTile, m - k0);
        __syncthreads();
        for (int t = threadIdx.x; t < kt; t += kNnThreads) {
            const float *q = known + (size_t)(k0 + t) * 3;
            sk[t] = make_float4(__ldg(q), __ldg(q + 1), __ldg(q + 2), 0.f);
        }
        __syncthreads();
        if (!live) continue;
#pragma unroll 4
        for (int k = 0; k < kt; ++k) {
            const float4 q = sk[k];
            const float d = sqdist(__fsub_rn(ux, q.x), __fsub_rn(uy, q.y), __fsub_rn(uz, q.z));
            if (d < b3) {                                        // rare after the first few candidates
                if (d < b1) { b3 = b2; i3 = i2; b2 = b1; i2 = i1; b1 = d; i1 = k0 + k; }
                else if (d < b2) { b3 = b2; i3 = i2; b2 = d; i2 = k0 + k; }
                else { b3 = d; i3 = k0 + k; }
            }
        }
    }
    if (!live) return;
    dist2[row * 3] = b1; dist2[row * 3 + 1] = b2; dist2[row * 3 + 2] = b3;
    idx[row * 3] = i1; idx[row * 3 + 1] = i2; idx[row * 3 + 2] = i3;
}

// out (B,C,N) = w0*points[i0] + w1*points[i1] + w2*points[i2], contracted as nvcc contracts the reference's
// expression (interpolate_gpu.cu:98; SASS: FMUL, FFMA, FFMA): fma(w2,p2, fma(w0,p0, w1*p1)).
template <bool kGrad>
__global__ void __launch_bounds__(kBtThreads)
three_interpolate_kernel(int C, int M, int N, const float *__restrict__ src, const int *__restrict__ idx,
                         const float *__restrict__ weight, float *__restrict__ dst) {
    const int b = blockIdx.z, p = blockIdx.x * kBtThreads + threadIdx.x;
    if (p >= N) return;
    const size_t row = ((size_t)b * N + p) * 3;
    const int i0 = __ldg(idx + row), i1 = __ldg(idx + row + 1), i2 = __ldg(idx + row + 2);
    const float w0 = __ldg(weight + row), w1 = __ldg(weight + row + 1), w2 = __ldg(weight + row + 2);
    constexpr int kCh = kGrad ? kBtChanGrad : kBtChan;
    const int c0 = blockIdx.y * kCh, c1 = min(C, c0 + kCh);
    for (int c = c0; c < c1; ++c) {
        const size_t plane = (size_t)b * C + c;
        if (kGrad) {                                             // src = grad_out (B,C,N), dst = grad_points (B,C,M)
            const float g = __ldg(src + plane * N + p);
            float *gp = dst + plane * M;
            atomicAdd(gp + i0, __fmul_rn(g, w0));
            atomicAdd(gp + i1, __fmul_rn(g, w1));
            atomicAdd(gp + i2, __fmul_rn(g, w2));
        } else {                                                 // src = points (B,C,M), dst = out (B,C,N)
            const float *pp = src + plane * M;
            dst[plane * N + p] = __fmaf_rn(w2, __ldg(pp + i2), __fmaf_rn(w0, __ldg(pp + i0), __fmul_rn(w1, __ldg(pp + i1))));
        }
    }
}

static int grid_ok(int b, int c, long long inner, const char *what) {
    if (b < 0 || c < 0 || inner < 0) { set_error("%s: negative size", what); return MGAR_EINVAL; }
    if (b > 65535 || div_up(c, kBtChanGrad) > 65535) { set_error("%s: b=%d / c=%d exceed the grid", what, b, c); return MGAR_EINVAL; }
    if (inner > 0x7fffffffLL) { set_error("%s: %lld positions exceed int32", what, inner); return MGAR_EINVAL; }
    return MGAR_OK;
}

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_batch_group_points(int b, int c, int n, int npoints, int nsample, const float *points, const int *idx,
                                       float *out, void *stream) {
    const long long P = (long long)npoints * nsample;
    if (int rc = grid_ok(b, c, P, "group_points(batch)")) return rc;
    if (b == 0 || c == 0 || P == 0) return MGAR_OK;
    MGAR_REQUIRE(points && idx && out && n > 0, "group_points(batch): null pointer or empty source");
    dim3 grid(div_up(P, kBtThreads), div_up(c, kBtChan), b);
    batch_gather_kernel<false><<<grid, kBtThreads, 0, (cudaStream_t)stream>>>(c, n, (int)P, points, idx, out);
    count_launch();
    return check_launch("group_points(batch)");
}

extern "C" int mgar_batch_group_points_grad(int b, int c, int n, int npoints, int nsample, const float *grad_out,
                                            const int *idx, float *grad_points, void *stream) {
    const long long P = (long long)npoints * nsample;
    if (int rc = grid_ok(b, c, P, "group_points_grad(batch)")) return rc;
    if (b == 0 || c == 0 || P == 0) return MGAR_OK;
    MGAR_REQUIRE(grad_out && idx && grad_points && n > 0, "group_points_grad(batch): null pointer or empty target");
    dim3 grid(div_up(P, kBtThreads), div_up(c, kBtChan), b);
    batch_gather_kernel<true><<<grid, kBtThreads, 0, (cudaStream_t)stream>>>(c, n, (int)P, grad_out, idx, grad_points);
    count_launch();
    return check_launch("group_points_grad(batch)");
}

extern "C" int mgar_three_nn(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                             void *stream) {
    if (int rc = grid_ok(b, 0, n, "three_nn")) return rc;
    MGAR_REQUIRE(m >= 0, "three_nn: negative size");
    if (b == 0 || n == 0) return MGAR_OK;
    MGAR_REQUIRE(unknown && dist2 && idx && (known || m == 0), "three_nn: null pointer");
    dim3 grid(div_up(n, kNnThreads), b);
    three_nn_kernel<<<grid, kNnThreads, 0, (cudaStream_t)stream>>>(n, m, unknown, known, dist2, idx);
    count_launch();
    return check_launch("three_nn");
}

extern "C" int mgar_three_interpolate(int b, int c, int m, int n, const float *points, const int *idx, const float *weight,
                                      float *out, void *stream) {
    if (int rc = grid_ok(b, c, n, "three_interpolate")) return rc;
    if (b == 0 || c == 0 || n == 0) return MGAR_OK;
    MGAR_REQUIRE(points && idx && weight && out && m > 0, "three_interpolate: null pointer or empty source");
    dim3 grid(div_up(n, kBtThreads), div_up(c, kBtChan), b);
    three_interpolate_kernel<false><<<grid, kBtThreads, 0, (cudaStream_t)stream>>>(c, m, n, points, idx, weight, out);
    count_launch();
    return check_launch("three_interpolate");
}

extern "C" int mgar_three_interpolate_grad(int b, int c, int n, int m, const float *grad_out, const int *idx,
                                           const float *weight, float *grad_points, void *stream) {
    if (int rc = grid_ok(b, c, n, "three_interpolate_grad")) return rc;
    if (b == 0 || c == 0 || n == 0) return MGAR_OK;
    MGAR_REQUIRE(grad_out && idx && weight && grad_points && m > 0, "three_interpolate_grad: null pointer or empty target");
    dim3 grid(div_up(n, kBtThreads), div_up(c, kBtChanGrad), b);
    three_interpolate_kernel<true><<<grid, kBtThreads, 0, (cudaStream_t)stream>>>(c, m, n, grad_out, idx, weight, grad_points);
    count_launch();
    return check_launch("three_interpolate_grad");
}
