// In-box tests: points_in_boxes_gpu (first hit), the (M,N) in-box matrix, and the host-buffer
// entry that replaces points_in_boxes_cpu.
//
// Reference: roiaware_pool3d/src/roiaware_pool3d_kernel.cu:313-336 (one thread per point,
// every thread re-reading all M*7 box floats from global and recomputing cosf/sinf per pair).
// Here: the box tile is pre-processed once per CTA into shared memory (trig, directed-rounded
// thresholds - see mgar_common.cuh), each thread owns 4 consecutive points read as three
// 128-bit loads and writes one 128-bit result; the kernel streams 12N + 4N bytes once (HBM-bound).
#include <math.h>


#include "mgar_common.cuh"

namespace mgar {

constexpr int kPibThreads = 256;
constexpr int kPibPts = 4;        // points per thread
constexpr int kBoxTile = 1024;    // boxes staged per pass (32 KB of shared memory)

// Load the 4 points owned by this thread.  Vector path when the frame base is 16B aligned.
__device__ __forceinline__ void load_pts4(const float *__restrict__ pts, long long p0, int n_valid,
                                          bool vec_ok, float (&x)[4], float (&y)[4], float (&z)[4]) {
    if (vec_ok && n_valid == 4) {
        const float4 *v = reinterpret_cast<const float4 *>(pts + p0 * 3);
        const float4 a = __ldg(v), b = __ldg(v + 1), c = __ldg(v + 2);
        x[0] = a.x; y[0] = a.y; z[0] = a.z;
        x[1] = a.w; y[1] = b.x; z[1] = b.y;
        x[2] = b.z; y[2] = b.w; z[2] = c.x;
        x[3] = c.y; y[3] = c.z; z[3] = c.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (j < n_valid) {
                x[j] = __ldg(pts + (p0 + j) * 3 + 0);
                y[j] = __ldg(pts + (p0 + j) * 3 + 1);
                z[j] = __ldg(pts + (p0 + j) * 3 + 2);
            } else {
                x[j] = y[j] = z[j] = nanf("");  // NaN never tests in-box
            }
        }
    }
}

__global__ void __launch_bounds__(kPibThreads)
points_in_boxes_first_kernel(const float *__restrict__ boxes, const float *__restrict__ pts,
                             int *__restrict__ box_idx, int M, int N, int vec_ok) {
    __shared__ BoxP sbox[kBoxTile];
    const int b = blockIdx.y;
    boxes += (size_t)b * M * 7;
    pts += (size_t)b * N * 3;
    box_idx += (size_t)b * N;

    const long long p0 = ((long long)blockIdx.x * kPibThreads + threadIdx.x) * kPibPts;
    const int n_valid = p0 >= N ? 0 : (N - p0 >= 4 ? 4 : (int)(N - p0));
    float x[4], y[4], z[4];
    load_pts4(pts, p0, n_valid, vec_ok != 0, x, y, z);

    int res[4] = {-1, -1, -1, -1};
    for (int k0 = 0; k0 < M; k0 += kBoxTile) {
        const int kt = min(kBoxTile, M - k0);
        __syncthreads();
        for (int k = threadIdx.x; k < kt; k += kPibThreads) sbox[k] = make_boxp<false>(boxes + (size_t)(k0 + k) * 7, nullptr);
        __syncthreads();
        for (int k = 0; k < kt; k++) {
            const BoxP bp = sbox[k];
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (res[j] < 0 && in_box<false>(x[j], y[j], z[j], bp)) res[j] = k0 + k;
        }
    }
    if (n_valid == 4 && (N % 4 == 0)) {
        *reinterpret_cast<int4 *>(box_idx + p0) = make_int4(res[0], res[1], res[2], res[3]);
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (j < n_valid) box_idx[p0 + j] = res[j];
    }
}

// (M,N) 0/1 matrix.  grid.x tiles points, grid.y tiles boxes (kMaskBoxes per CTA).
constexpr int kMaskBoxes = 8;

template <bool kCpuSemantics>
__global__ void __launch_bounds__(kPibThreads)
in_box_mask_kernel(const float *__restrict__ boxes, const float *__restrict__ pts,
                   const float *__restrict__ cs, int *__restrict__ mask, int M, int N, int vec_ok) {
    __shared__ BoxP sbox[kMaskBoxes];
    const int k0 = blockIdx.y * kMaskBoxes;
    const int kt = min(kMaskBoxes, M - k0);
    if (threadIdx.x < kt)
        sbox[threadIdx.x] = make_boxp<kCpuSemantics>(boxes + (size_t)(k0 + threadIdx.x) * 7,
                                                     kCpuSemantics ? cs + (size_t)(k0 + threadIdx.x) * 2 : nullptr);
    const long long p0 = ((long long)blockIdx.x * kPibThreads + threadIdx.x) * kPibPts;
    const int n_valid = p0 >= N ? 0 : (N - p0 >= 4 ? 4 : (int)(N - p0));
    float x[4], y[4], z[4];
    load_pts4(pts, p0, n_valid, vec_ok != 0, x, y, z);
    __syncthreads();
    for (int k = 0; k < kt; k++) {
        const BoxP bp = sbox[k];
        int r[4];
#pragma unroll
        for (int j = 0; j < 4; j++) r[j] = in_box<kCpuSemantics>(x[j], y[j], z[j], bp) ? 1 : 0;
        int *row = mask + (size_t)(k0 + k) * N;
        if (n_valid == 4 && (N % 4 == 0)) {
            __stcs(reinterpret_cast<int4 *>(row + p0), make_int4(r[0], r[1], r[2], r[3]));
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (j < n_valid) row[p0 + j] = r[j];
        }
    }
}

static bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_points_in_boxes(const float *boxes, const float *pts, int *box_idx, int B, int M,
                                    int N, void *stream) {
    MGAR_REQUIRE(B >= 0 && M >= 0 && N >= 0, "points_in_boxes: negative size B=%d M=%d N=%d", B, M, N);
    if (B == 0 || N == 0) return MGAR_OK;
    MGAR_REQUIRE(pts && box_idx && (boxes || M == 0), "points_in_boxes: null pointer");
    MGAR_REQUIRE(B <= 65535, "points_in_boxes: B=%d exceeds grid.y limit", B);
    const int vec_ok = aligned16(pts) && aligned16(box_idx) && (N % 4 == 0);
    dim3 grid(div_up(N, kPibThreads * kPibPts), B);
    points_in_boxes_first_kernel<<<grid, kPibThreads, 0, (cudaStream_t)stream>>>(boxes, pts, box_idx, M, N, vec_ok);
    count_launch();
    return check_launch("points_in_boxes");
}

extern "C" int mgar_in_box_mask(const float *boxes, const float *pts, const float *cs, int *mask, int M,
                                int N, int cpu_semantics, void *stream) {
    MGAR_REQUIRE(M >= 0 && N >= 0, "in_box_mask: negative size");
    if (M == 0 || N == 0) return MGAR_OK;
    MGAR_REQUIRE(boxes && pts && mask, "in_box_mask: null pointer");
    MGAR_REQUIRE(!cpu_semantics || cs, "in_box_mask: cpu_semantics needs host-computed cos/sin");
    MGAR_REQUIRE(div_up(M, kMaskBoxes) <= 65535, "in_box_mask: too many boxes");
    const int vec_ok = aligned16(pts) && aligned16(mask) && (N % 4 == 0);
    dim3 grid(div_up(N, kPibThreads * kPibPts), div_up(M, kMaskBoxes));
    if (cpu_semantics)
        in_box_mask_kernel<true><<<grid, kPibThreads, 0, (cudaStream_t)stream>>>(boxes, pts, cs, mask, M, N, vec_ok);
    else
        in_box_mask_kernel<false><<<grid, kPibThreads, 0, (cudaStream_t)stream>>>(boxes, pts, cs, mask, M, N, vec_ok);
    count_launch();
    return check_launch("in_box_mask");
}

// device staging of the host-buffer entry: boxes | cos, sin | points | (M,N) mask, each rounded up to 256 bytes
extern "C" size_t mgar_points_in_boxes_mask_host_workspace_bytes(int M, int N) {
    if (M <= 0 || N <= 0) return 0;
    return up256((size_t)M * 28) + up256((size_t)M * 8) + up256((size_t)N * 12) + up256((size_t)M * N * 4);
}

extern "C" int mgar_points_in_boxes_mask_host(const float *boxes_host, const float *pts_host,
                                              int *mask_host, int M, int N, void *workspace, size_t workspace_bytes) {
    MGAR_REQUIRE(M >= 0 && N >= 0, "points_in_boxes_mask_host: negative size");
    if (M == 0 || N == 0) return MGAR_OK;
    MGAR_REQUIRE(boxes_host && pts_host && mask_host, "points_in_boxes_mask_host: null pointer");
    const size_t o_box = 0, o_cs = up256((size_t)M * 28), o_pts = o_cs + up256((size_t)M * 8);
    const size_t o_mask = o_pts + up256((size_t)N * 12), total = o_mask + up256((size_t)M * N * 4);
    if (!workspace || workspace_bytes < total) {      // the caller owns the device staging (current device): no hidden allocation
        set_error("points_in_boxes_mask_host: needs a DEVICE workspace of mgar_points_in_boxes_mask_host_workspace_bytes(M, N) "
                  "= %zu bytes, got %zu", total, workspace ? workspace_bytes : (size_t)0);
        return MGAR_ENOMEM;
    }
    char *d = static_cast<char *>(workspace);
    // The only host arithmetic: the M box headings go through the host libm so the decision is
    // bit-identical to the reference's CPU build (roiaware_pool3d.cpp:121-125).
    float *cs = new float[(size_t)M * 2];
    for (int k = 0; k < M; k++) {
        const float rz = boxes_host[(size_t)k * 7 + 6];
        cs[2 * k] = cosf(-rz);
        cs[2 * k + 1] = sinf(-rz);
    }
    cudaStream_t st = cudaStreamPerThread;
    cudaError_t e = cudaMemcpyAsync(d + o_box, boxes_host, (size_t)M * 28, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + o_cs, cs, (size_t)M * 8, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + o_pts, pts_host, (size_t)N * 12, cudaMemcpyHostToDevice, st);
    int rc = MGAR_OK;
    if (e == cudaSuccess)
        rc = mgar_in_box_mask((const float *)(d + o_box), (const float *)(d + o_pts), (const float *)(d + o_cs),
                              (int *)(d + o_mask), M, N, 1, st);
    if (e == cudaSuccess && rc == MGAR_OK)
        e = cudaMemcpyAsync(mask_host, d + o_mask, (size_t)M * N * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && rc == MGAR_OK) e = cudaStreamSynchronize(st);
    delete[] cs;
    if (rc != MGAR_OK) return rc;
    if (e != cudaSuccess) {
        set_error("points_in_boxes_mask_host: %s", cudaGetErrorString(e));
        return MGAR_ECUDA;
    }
    return MGAR_OK;
}
