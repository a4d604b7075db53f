// Stack-layout PointNet++ ops: ball query and grouping (+ gradient).
//
// Reference: pointnet2/pointnet2_stack/src/ball_query_gpu.cu:16-66 (one THREAD per centroid
// walking its cloud serially from global memory, K scattered stores on the first hit, an O(B)
// loop per thread to find its batch) and group_points_gpu.cu:15-102 (one thread per element,
// the same O(B) loop each).
//
// Here: a tiny scan kernel turns the ragged counts into prefix arrays once; ball query gives a
// WARP to each centroid and a CTA to a tile of 64 centroids of one cloud - the cloud is staged in
// shared memory in 4096-point chunks (coalesced, read once per CTA), lanes test 32 points per
// step and append in index order with ballot/popc, and each result row leaves as one coalesced
// 4K-byte store.  Distances keep the reference's exact FMUL/FFMA/FFMA order and strict '<'.
#include "mgar_common.cuh"
#include "stack_prefix.cuh"

namespace mgar {

constexpr int kBqTile = 64;       // centroids per CTA
constexpr int kBqWarps = 8;
constexpr int kBqPerWarp = kBqTile / kBqWarps;
constexpr int kBqChunk = 4096;    // cloud points staged per pass (48 KB)

// prefix[0..B] for two count arrays + tile prefix for the first one (tiles of `tile` rows).
// Single CTA; B is small (persons or frames).
__global__ void __launch_bounds__(1024)
stack_prefix_kernel(const int *__restrict__ cnt_a, const int *__restrict__ cnt_b, int B, int tile,
                    int *__restrict__ pre_a, int *__restrict__ pre_b, int *__restrict__ pre_tile) {
    __shared__ int s_w[3][32];
    __shared__ int s_carry[3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 3) s_carry[tid] = 0;
    __syncthreads();
    for (int base = 0; base < B; base += 1024) {
        const int k = base + tid;
        int v[3];
        v[0] = k < B ? cnt_a[k] : 0;
        v[1] = (k < B && cnt_b) ? cnt_b[k] : 0;
        v[2] = (v[0] + tile - 1) / tile;
        int inc[3];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            int x = v[q];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, x, d);
                if (lane >= d) x += t;
            }
            inc[q] = x;
            if (lane == 31) s_w[q][warp] = x;
        }
        __syncthreads();
        if (warp == 0) {
#pragma unroll
            for (int q = 0; q < 3; q++) {
                int x = s_w[q][lane];
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, x, d);
                    if (lane >= d) x += t;
                }
                s_w[q][lane] = x;  // inclusive over warps
            }
        }
        __syncthreads();
        int *outs[3] = {pre_a, pre_b, pre_tile};
#pragma unroll
        for (int q = 0; q < 3; q++) {
            const int excl = s_carry[q] + (warp ? s_w[q][warp - 1] : 0) + inc[q] - v[q];
            if (k < B && outs[q]) outs[q][k] = excl;
        }
        __syncthreads();
        if (tid == 0)
            for (int q = 0; q < 3; q++) s_carry[q] += s_w[q][31];
        __syncthreads();
    }
    if (tid == 0) {
        if (pre_a) pre_a[B] = s_carry[0];
        if (pre_b) pre_b[B] = s_carry[1];
        if (pre_tile) pre_tile[B] = s_carry[2];
    }
}

// largest b in [0,B) with pre[b] <= v  (pre is a non-decreasing exclusive prefix, pre[B] = total).
// Matches the reference's batch search for every v < total (empty batches are skipped).
__device__ __forceinline__ int find_batch(const int *__restrict__ pre, int B, int v) {
    int lo = 0, hi = B;  // invariant: pre[lo] <= v < pre[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(pre + mid) <= v) lo = mid; else hi = mid;
    }
    return lo;
}

// uniform != 0: every cloud has n_u points / m_u centroids and the prefix arrays are not read.
__global__ void __launch_bounds__(kBqWarps * 32)
ball_query_kernel(int B, float radius, int K, const float *__restrict__ new_xyz, const float *__restrict__ xyz,
                  const int *__restrict__ pre_new, const int *__restrict__ pre_xyz,
                  const int *__restrict__ pre_tile, int uniform, int n_u, int m_u, int *__restrict__ idx) {
    extern __shared__ float smem_f[];
    float *sx = smem_f;                                              // [kBqChunk][3]
    int *sidx = reinterpret_cast<int *>(smem_f + kBqChunk * 3);      // [kBqTile][K]

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int b, c_begin, c_end, n;
    long long p_start;
    if (uniform) {
        const int tiles_per = (m_u + kBqTile - 1) / kBqTile;
        b = blockIdx.x / tiles_per;
        const int t = blockIdx.x - b * tiles_per;
        c_begin = b * m_u + t * kBqTile;
        c_end = min(c_begin + kBqTile, (b + 1) * m_u);
        n = n_u;
        p_start = (long long)b * n_u;
    } else {
        if ((int)blockIdx.x >= __ldg(pre_tile + B)) return;
        b = find_batch(pre_tile, B, blockIdx.x);
        const int t = blockIdx.x - __ldg(pre_tile + b);
        c_begin = __ldg(pre_new + b) + t * kBqTile;
        c_end = min(c_begin + kBqTile, __ldg(pre_new + b + 1));
        n = __ldg(pre_xyz + b + 1) - __ldg(pre_xyz + b);
        p_start = __ldg(pre_xyz + b);
    }
    xyz += p_start * 3;
    const float r2 = __fmul_rn(radius, radius);

    float qx[kBqPerWarp], qy[kBqPerWarp], qz[kBqPerWarp];
    int cnt[kBqPerWarp], kfirst[kBqPerWarp];
#pragma unroll
    for (int c = 0; c < kBqPerWarp; c++) {
        const int ci = c_begin + warp * kBqPerWarp + c;
        const bool ok = ci < c_end;
        qx[c] = ok ? __ldg(new_xyz + (size_t)ci * 3 + 0) : 0.f;
        qy[c] = ok ? __ldg(new_xyz + (size_t)ci * 3 + 1) : 0.f;
        qz[c] = ok ? __ldg(new_xyz + (size_t)ci * 3 + 2) : 0.f;
        cnt[c] = ok ? 0 : K;  // out-of-tile slots count as finished
        kfirst[c] = 0;
    }

    for (int k0 = 0; k0 < n; k0 += kBqChunk) {
        const int kn = min(kBqChunk, n - k0);
        __syncthreads();
        for (int i = threadIdx.x; i < kn * 3; i += kBqWarps * 32) sx[i] = __ldg(xyz + (size_t)k0 * 3 + i);
        __syncthreads();
        bool all_done = true;
#pragma unroll
        for (int c = 0; c < kBqPerWarp; c++) {
            int *row = sidx + (warp * kBqPerWarp + c) * K;
            for (int base = 0; base < kn && cnt[c] < K; base += 32) {
                const int k = base + lane;
                bool hit = false;
                if (k < kn) {
                    const float d2 = sqdist(__fsub_rn(qx[c], sx[k * 3 + 0]), __fsub_rn(qy[c], sx[k * 3 + 1]),
                                            __fsub_rn(qz[c], sx[k * 3 + 2]));
                    hit = d2 < r2;
                }
                const unsigned msk = __ballot_sync(0xffffffffu, hit);
                if (msk == 0) continue;
                if (cnt[c] == 0) kfirst[c] = k0 + base + __ffs(msk) - 1;
                const int pos = cnt[c] + __popc(msk & ((1u << lane) - 1u));
                if (hit && pos < K) row[pos] = k0 + k;
                cnt[c] += __popc(msk);
            }
            all_done &= cnt[c] >= K;
        }
        if (__syncthreads_and(all_done)) break;
    }
    __syncwarp();
#pragma unroll
    for (int c = 0; c < kBqPerWarp; c++) {
        const int ci = c_begin + warp * kBqPerWarp + c;
        if (ci >= c_end) continue;
        const int *row = sidx + (warp * kBqPerWarp + c) * K;
        int *out = idx + (size_t)ci * K;
        if (cnt[c] == 0) {
            if (lane == 0) out[0] = -1;  // other slots keep the caller's fill (ball_query_gpu.cu:65)
        } else {
            for (int l = lane; l < K; l += 32) out[l] = l < cnt[c] ? row[l] : kfirst[c];
        }
    }
}

__global__ void __launch_bounds__(256)
group_points_kernel(int B, long long total, int C, int K, const float *__restrict__ features,
                    const int *__restrict__ pre_feat, const int *__restrict__ pre_idx, const int *__restrict__ idx,
                    float *__restrict__ out) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= total) return;
    const int s = (int)(i % K);
    const int c = (int)((i / K) % C);
    const int pt = (int)(i / K / C);
    const int b = find_batch(pre_idx, B, pt);
    const long long row = (long long)__ldg(pre_feat + b) + __ldg(idx + (size_t)pt * K + s);
    out[i] = __ldg(features + row * C + c);
}

__global__ void __launch_bounds__(256)
group_points_grad_kernel(int B, long long total, int C, int K, const float *__restrict__ grad_out,
                         const int *__restrict__ pre_feat, const int *__restrict__ pre_idx,
                         const int *__restrict__ idx, float *__restrict__ grad_features) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= total) return;
    const int s = (int)(i % K);
    const int c = (int)((i / K) % C);
    const int pt = (int)(i / K / C);
    const int b = find_batch(pre_idx, B, pt);
    const long long row = (long long)__ldg(pre_feat + b) + __ldg(idx + (size_t)pt * K + s);
    atomicAdd(grad_features + row * C + c, __ldg(grad_out + i));
}

int stack_prefix(const int *cnt_a, const int *cnt_b, int B, int tile, int *pre_a, int *pre_b, int *pre_tile,
                 cudaStream_t st) {
    stack_prefix_kernel<<<1, 1024, 0, st>>>(cnt_a, cnt_b, B, tile, pre_a, pre_b, pre_tile);
    count_launch();
    return check_launch("stack_prefix");
}

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_ball_query_stack(int B, int M, float radius, int nsample, const float *new_xyz,
                                     const int *new_xyz_batch_cnt, const float *xyz, const int *xyz_batch_cnt,
                                     int *idx, void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0 && nsample >= 1, "ball_query: bad size B=%d M=%d nsample=%d", B, M, nsample);
    if (M == 0) return MGAR_OK;
    MGAR_REQUIRE(new_xyz && new_xyz_batch_cnt && xyz && xyz_batch_cnt && idx, "ball_query: null pointer");
    const size_t smem = (size_t)kBqChunk * 12 + (size_t)kBqTile * nsample * sizeof(int);
    MGAR_REQUIRE(smem <= 200 * 1024, "ball_query: nsample=%d too large", nsample);
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (int rc = ws.alloc(B, st)) return rc;
    if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, kBqTile, ws.a, ws.b, ws.t, st)) return rc;
    MGAR_CUDA(cudaFuncSetAttribute(ball_query_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = div_up(M, kBqTile) + B;  // upper bound of sum_b ceil(M_b / tile)
    ball_query_kernel<<<grid, kBqWarps * 32, smem, st>>>(B, radius, nsample, new_xyz, xyz, ws.a, ws.b, ws.t, 0, 0, 0, idx);
    count_launch();
    return check_launch("ball_query");
}

extern "C" int mgar_group_points_stack(int B, int M, int C, int nsample, const float *features,
                                       const int *features_batch_cnt, const int *idx, const int *idx_batch_cnt,
                                       float *out, void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0 && C >= 0 && nsample >= 0, "group_points: bad size");
    const long long total = (long long)M * C * nsample;
    if (total == 0) return MGAR_OK;
    MGAR_REQUIRE(features && features_batch_cnt && idx && idx_batch_cnt && out, "group_points: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (int rc = ws.alloc(B, st)) return rc;
    if (int rc = stack_prefix(idx_batch_cnt, features_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    group_points_kernel<<<div_up(total, 256), 256, 0, st>>>(B, total, C, nsample, features, ws.b, ws.a, idx, out);
    count_launch();
    return check_launch("group_points");
}

extern "C" int mgar_group_points_grad_stack(int B, int M, int C, int N, int nsample, const float *grad_out,
                                            const int *idx, const int *idx_batch_cnt,
                                            const int *features_batch_cnt, float *grad_features, void *stream) {
    (void)N;
    MGAR_REQUIRE(B >= 1 && M >= 0 && C >= 0 && nsample >= 0, "group_points_grad: bad size");
    const long long total = (long long)M * C * nsample;
    if (total == 0) return MGAR_OK;
    MGAR_REQUIRE(grad_out && features_batch_cnt && idx && idx_batch_cnt && grad_features, "group_points_grad: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (int rc = ws.alloc(B, st)) return rc;
    if (int rc = stack_prefix(idx_batch_cnt, features_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    group_points_grad_kernel<<<div_up(total, 256), 256, 0, st>>>(B, total, C, nsample, grad_out, ws.b, ws.a, idx, grad_features);
    count_launch();
    return check_launch("group_points_grad");
}
