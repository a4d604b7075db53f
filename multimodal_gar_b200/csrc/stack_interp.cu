// Stack-layout three-nearest-neighbour search and feature interpolation (PointNet++ feature propagation on ragged
// batches): pointnet2_stack_cuda.three_nn_wrapper / three_interpolate_wrapper / three_interpolate_grad_wrapper.
//
// Reference: pointnet2_stack/src/interpolate_gpu.cu:16-74 (three_nn: one thread per unknown point, an O(B) batch
// search per thread, then a serial walk over its cloud's known points straight from global memory, bests kept as
// doubles), :107-126 (interpolate: one thread per OUTPUT ELEMENT with the channel in blockIdx.y, so a warp reads 32
// feature rows C floats apart and re-reads every index / weight C times), :150-170 (grad: same shape, 3 atomics).
// Here: three_nn = one CTA per tile of 128 unknown points of ONE cloud (tile -> cloud through the prefix of
// per-cloud tile counts), known points staged through shared memory as float4 (one broadcast LDS.128 per candidate
// per warp); interpolate / grad = channel is the fastest thread index, so the three gathered feature rows are read
// as contiguous 128-byte segments and the output row is written coalesced (float4 when C % 4 == 0).
#include "stack_prefix.cuh"

namespace mgar {

constexpr int kSnThreads = 128;   // unknown points per CTA
constexpr int kSnTile = 1024;     // known points per shared-memory chunk (16 KB)

// Strict '<' against the running bests while walking k upward: ties keep the lowest index (reference :55-68; its
// double bests only ever hold float values or 1e40, so float comparisons against +inf decide identically, and
// (float)1e40 = +inf is what the reference stores when a cloud has fewer than three known points).
__global__ void __launch_bounds__(kSnThreads)
three_nn_stack_kernel(int B, const float *__restrict__ unknown, const float *__restrict__ known,
                      const int *__restrict__ pre_unknown, const int *__restrict__ pre_known,
                      const int *__restrict__ pre_tile, float *__restrict__ dist2, int *__restrict__ idx) {
    __shared__ float4 sk[kSnTile];
    if ((int)blockIdx.x >= __ldg(pre_tile + B)) return;
    const int b = find_batch(pre_tile, B, blockIdx.x);
    const int p = __ldg(pre_unknown + b) + ((int)blockIdx.x - __ldg(pre_tile + b)) * kSnThreads + threadIdx.x;
    const bool live = p < __ldg(pre_unknown + b + 1);
    const int k_start = __ldg(pre_known + b), m = __ldg(pre_known + b + 1) - k_start;
    const size_t row = live ? (size_t)p : 0;
    const float ux = __ldg(unknown + row * 3), uy = __ldg(unknown + row * 3 + 1), uz = __ldg(unknown + row * 3 + 2);
    known += (size_t)k_start * 3;
    float b1 = INFINITY, b2 = INFINITY, b3 = INFINITY;
    int i1 = 0, i2 = 0, i3 = 0;
    for (int k0 = 0; k0 < m; k0 += kSnTile) {
        const int kt = min(kSnTile, m - k0);
        __syncthreads();
        for (int t = threadIdx.x; t < kt; t += kSnThreads) {
            const float *q = known + (size_t)(k0 + t) * 3;
            sk[t] = make_float4(__ldg(q), __ldg(q + 1), __ldg(q + 2), 0.f);
        }
        __syncthreads();
        if (!live) continue;
#pragma unroll 4
        for (int k = 0; k < kt; ++k) {
            const float4 q = sk[k];
            const float d = sqdist(__fsub_rn(ux, q.x), __fsub_rn(uy, q.y), __fsub_rn(uz, q.z));
            if (d < b3) {
                if (d < b1) { b3 = b2; i3 = i2; b2 = b1; i2 = i1; b1 = d; i1 = k0 + k; }
                else if (d < b2) { b3 = b2; i3 = i2; b2 = d; i2 = k0 + k; }
                else { b3 = d; i3 = k0 + k; }
            }
        }
    }
    if (!live) return;
    dist2[row * 3] = b1; dist2[row * 3 + 1] = b2; dist2[row * 3 + 2] = b3;
    idx[row * 3] = i1 + k_start; idx[row * 3 + 1] = i2 + k_start; idx[row * 3 + 2] = i3 + k_start;   // global rows (:70-72)
}

// out (N,C) = w0*f[i0] + w1*f[i1] + w2*f[i2] contracted as nvcc contracts the reference's expression (:123-125;
// SASS of the reference build: FMUL w1*f1, FFMA w0*f0, FFMA w2*f2).
__device__ __forceinline__ float interp3(float w0, float f0, float w1, float f1, float w2, float f2) {
    return __fmaf_rn(w2, f2, __fmaf_rn(w0, f0, __fmul_rn(w1, f1)));
}

template <int kVec>
__global__ void __launch_bounds__(256)
three_interpolate_stack_kernel(long long total, int Cv, const float *__restrict__ features, const int *__restrict__ idx,
                               const float *__restrict__ weight, float *__restrict__ out) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;           // (point, channel vector), channel fastest
    if (i >= total) return;
    const long long pt = i / Cv;
    const int cv = (int)(i - pt * Cv);
    const int i0 = __ldg(idx + pt * 3), i1 = __ldg(idx + pt * 3 + 1), i2 = __ldg(idx + pt * 3 + 2);
    const float w0 = __ldg(weight + pt * 3), w1 = __ldg(weight + pt * 3 + 1), w2 = __ldg(weight + pt * 3 + 2);
    if (kVec == 4) {
        const float4 *f = reinterpret_cast<const float4 *>(features);
        const float4 a = __ldg(f + (size_t)i0 * Cv + cv), b = __ldg(f + (size_t)i1 * Cv + cv), c = __ldg(f + (size_t)i2 * Cv + cv);
        reinterpret_cast<float4 *>(out)[i] = make_float4(interp3(w0, a.x, w1, b.x, w2, c.x), interp3(w0, a.y, w1, b.y, w2, c.y),
                                                         interp3(w0, a.z, w1, b.z, w2, c.z), interp3(w0, a.w, w1, b.w, w2, c.w));
    } else {
        out[i] = interp3(w0, __ldg(features + (size_t)i0 * Cv + cv), w1, __ldg(features + (size_t)i1 * Cv + cv), w2,
                         __ldg(features + (size_t)i2 * Cv + cv));
    }
}

// grad_features (M,C) += grad_out (N,C) * w scattered to the three source rows (reference :165-167, FMUL then RED).
__global__ void __launch_bounds__(256)
three_interpolate_grad_stack_kernel(long long total, int C, const float *__restrict__ grad_out, const int *__restrict__ idx,
                                    const float *__restrict__ weight, float *__restrict__ grad_features) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= total) return;
    const long long pt = i / C;
    const int c = (int)(i - pt * C);
    const float g = __ldg(grad_out + i);
#pragma unroll
    for (int j = 0; j < 3; ++j)
        atomicAdd(grad_features + (size_t)__ldg(idx + pt * 3 + j) * C + c, __fmul_rn(g, __ldg(weight + pt * 3 + j)));
}

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_three_nn_stack(int B, int N, int M, const float *unknown, const int *unknown_batch_cnt,
                                   const float *known, const int *known_batch_cnt, float *dist2, int *idx, void *workspace,
                                   size_t workspace_bytes, void *stream) {
    MGAR_REQUIRE(B >= 1 && N >= 0 && M >= 0, "three_nn_stack: bad size B=%d N=%d M=%d", B, N, M);
    if (N == 0) return MGAR_OK;
    MGAR_REQUIRE(unknown && unknown_batch_cnt && known_batch_cnt && dist2 && idx && (known || M == 0),
                 "three_nn_stack: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (int rc = ws.bind(workspace, workspace_bytes, B, "three_nn_stack")) return rc;
    if (int rc = stack_prefix(unknown_batch_cnt, known_batch_cnt, B, kSnThreads, ws.a, ws.b, ws.t, st)) return rc;
    const int grid = div_up(N, kSnThreads) + B;      // upper bound of sum(ceil(n_b / tile)); surplus CTAs exit at once
    three_nn_stack_kernel<<<grid, kSnThreads, 0, st>>>(B, unknown, known, ws.a, ws.b, ws.t, dist2, idx);
    count_launch();
    return check_launch("three_nn_stack");
}

extern "C" int mgar_three_interpolate_stack(int N, int C, const float *features, const int *idx, const float *weight,
                                            float *out, void *stream) {
    MGAR_REQUIRE(N >= 0 && C >= 0, "three_interpolate_stack: negative size");
    if ((long long)N * C == 0) return MGAR_OK;
    MGAR_REQUIRE(features && idx && weight && out, "three_interpolate_stack: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = C % 4 == 0 && ((reinterpret_cast<uintptr_t>(features) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    const long long total = (long long)N * (vec ? C / 4 : C);
    MGAR_REQUIRE(div_up(total, 256) > 0, "three_interpolate_stack: %lld elements exceed the grid", total);
    if (vec) three_interpolate_stack_kernel<4><<<div_up(total, 256), 256, 0, st>>>(total, C / 4, features, idx, weight, out);
    else three_interpolate_stack_kernel<1><<<div_up(total, 256), 256, 0, st>>>(total, C, features, idx, weight, out);
    count_launch();
    return check_launch("three_interpolate_stack");
}

extern "C" int mgar_three_interpolate_grad_stack(int N, int C, const float *grad_out, const int *idx, const float *weight,
                                                 float *grad_features, void *stream) {
    MGAR_REQUIRE(N >= 0 && C >= 0, "three_interpolate_grad_stack: negative size");
    const long long total = (long long)N * C;
    if (total == 0) return MGAR_OK;
    MGAR_REQUIRE(grad_out && idx && weight && grad_features, "three_interpolate_grad_stack: null pointer");
    MGAR_REQUIRE(div_up(total, 256) > 0, "three_interpolate_grad_stack: %lld elements exceed the grid", total);
    three_interpolate_grad_stack_kernel<<<div_up(total, 256), 256, 0, (cudaStream_t)stream>>>(total, C, grad_out, idx, weight,
                                                                                               grad_features);
    count_launch();
    return check_launch("three_interpolate_grad_stack");
}
