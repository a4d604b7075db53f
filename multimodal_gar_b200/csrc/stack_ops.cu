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

constexpr int kBqTile = 128;      // centroids (= threads) per CTA
constexpr int kBqChunk = 1024;    // cloud points staged per pass (16 KB as float4)

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

// One THREAD per centroid, one CTA per tile of 128 centroids of one cloud.  The cloud is staged in
// shared memory as float4 in 1024-point chunks; every lane of a warp reads the SAME point (one
// broadcast LDS.128 serves 32 distance tests), hits are appended in index order to the thread's own
// shared-memory row, and a warp stops scanning once all of its 32 centroids are full.  Rows leave as
// coalesced 4K-byte stores.  uniform != 0: every cloud has n_u points / m_u centroids and the
// prefix arrays are not read.  fill == 1: rows without a neighbour are written as (-1, 0, 0, ...)
// instead of relying on the caller's zero fill; fill == 2: such rows are not touched at all (pointnet2_batch).
__global__ void __launch_bounds__(kBqTile)
ball_query_kernel(int B, float radius, int K, const float *__restrict__ new_xyz, const float *__restrict__ xyz, int xs,
                  const int *__restrict__ pre_new, const int *__restrict__ pre_xyz,
                  const int *__restrict__ pre_tile, int uniform, int n_u, int m_u, int *__restrict__ idx, int fill) {
    extern __shared__ float4 smem_q[];
    float4 *sp = smem_q;                                             // [kBqChunk]
    int *rows = reinterpret_cast<int *>(smem_q + kBqChunk);          // [kBqTile][K + 1]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
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
    xyz += p_start * xs;
    const float r2 = __fmul_rn(radius, radius);
    const bool vec4 = (xs == 4) && ((reinterpret_cast<uintptr_t>(xyz) & 15) == 0);

    const int ci = c_begin + tid;
    const bool ok = ci < c_end;
    const float qx = ok ? __ldg(new_xyz + (size_t)ci * 3 + 0) : 0.f;
    const float qy = ok ? __ldg(new_xyz + (size_t)ci * 3 + 1) : 0.f;
    const float qz = ok ? __ldg(new_xyz + (size_t)ci * 3 + 2) : 0.f;
    int cnt = ok ? 0 : K;                                            // out-of-tile threads count as finished
    int *myrow = rows + tid * (K + 1);

    for (int k0 = 0; k0 < n; k0 += kBqChunk) {
        const int kn = min(kBqChunk, n - k0);
        __syncthreads();
        if (vec4) {
            for (int i = tid; i < kn; i += kBqTile) sp[i] = __ldg(reinterpret_cast<const float4 *>(xyz) + k0 + i);
        } else {
            for (int i = tid; i < kn; i += kBqTile) {
                const float *p = xyz + (size_t)(k0 + i) * xs;
                sp[i] = make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), 0.f);
            }
        }
        __syncthreads();
        for (int k = 0; k < kn; k += 4) {
            if (__all_sync(0xffffffffu, cnt >= K)) break;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 p = sp[min(k + u, kn - 1)];
                const float d2 = sqdist(__fsub_rn(qx, p.x), __fsub_rn(qy, p.y), __fsub_rn(qz, p.z));
                if (d2 < r2 && k + u < kn && cnt < K) {
                    myrow[cnt] = k0 + k + u;
                    cnt++;
                }
            }
        }
        if (__syncthreads_and(cnt >= K)) break;
    }
    __syncwarp();
    for (int r = 0; r < 32; r++) {
        const int row = warp * 32 + r;
        if (c_begin + row >= c_end) break;
        const int cnt_r = __shfl_sync(0xffffffffu, cnt, r);
        const int *src = rows + row * (K + 1);
        int *out = idx + (size_t)(c_begin + row) * K;
        if (cnt_r == 0) {
            if (fill == 1) { for (int l = lane; l < K; l += 32) out[l] = l == 0 ? -1 : 0; }
            else if (fill == 0 && lane == 0) out[0] = -1;   // other slots keep the caller's fill (ball_query_gpu.cu:65)
            // fill == 2: pointnet2_batch semantics, the whole row keeps the caller's fill (no -1 marker)
        } else {
            const int first = src[0];
            for (int l = lane; l < K; l += 32) out[l] = l < cnt_r ? src[l] : first;
        }
    }
}

// Uniform layout, clouds of at most 512 points: one CTA per cloud, a WARP per centroid, and the cloud lives in
// REGISTERS (lane l owns points l, l+32, ...: PPL x 3 floats) for all the centroids the warp serves - the inner
// loop touches no memory at all.  A step tests 64 points (two per lane), appends the hits in index order with
// ballot/popc into the warp's shared-memory row and stops as soon as THIS centroid has nsample neighbours; in
// the thread-per-centroid kernel above a warp scans until the slowest of its 32 centroids is full, which on
// pooled person clouds (sparse extremities picked by FPS) is nearly always the whole cloud.  The row leaves as
// one coalesced store (lane = slot).
constexpr int kBqRegWarps = 8;
template <int PPL>
__global__ void __launch_bounds__(kBqRegWarps * 32)
ball_query_reg_kernel(float radius, int K, const float *__restrict__ new_xyz, const float *__restrict__ xyz, int xs,
                      int n_u, int m_u, int *__restrict__ idx, int fill) {
    extern __shared__ int bq_rows[];                                 // [kBqRegWarps][K] rows, then [m_u * 3] centroids
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.x;
    float *sq = reinterpret_cast<float *>(bq_rows + kBqRegWarps * K);
    for (int i = threadIdx.x; i < m_u * 3; i += kBqRegWarps * 32) sq[i] = __ldg(new_xyz + (size_t)b * m_u * 3 + i);
    __syncthreads();
    xyz += (size_t)b * n_u * xs;
    const bool vec4 = (xs == 4) && ((reinterpret_cast<uintptr_t>(xyz) & 15) == 0);
    float px[PPL], py[PPL], pz[PPL];
#pragma unroll
    for (int s = 0; s < PPL; s++) {
        const int k = s * 32 + lane;
        px[s] = py[s] = pz[s] = __int_as_float(0x7f800000);          // +inf: never inside a ball
        if (k < n_u) {
            if (vec4) {
                const float4 p = __ldg(reinterpret_cast<const float4 *>(xyz) + k);
                px[s] = p.x; py[s] = p.y; pz[s] = p.z;
            } else {
                const float *p = xyz + (size_t)k * xs;
                px[s] = __ldg(p); py[s] = __ldg(p + 1); pz[s] = __ldg(p + 2);
            }
        }
    }
    const float r2 = __fmul_rn(radius, radius);
    const unsigned lt = (1u << lane) - 1u;
    int *row = bq_rows + warp * K;
    for (int c = warp; c < m_u; c += kBqRegWarps) {
        const size_t ci = (size_t)b * m_u + c;
        const float qx = sq[c * 3 + 0], qy = sq[c * 3 + 1], qz = sq[c * 3 + 2];
        int cnt = 0;
#pragma unroll
        for (int s = 0; s < PPL; s += 2) {
            const bool h0 = sqdist(__fsub_rn(qx, px[s]), __fsub_rn(qy, py[s]), __fsub_rn(qz, pz[s])) < r2;
            const bool h1 = (s + 1 < PPL) &&
                            sqdist(__fsub_rn(qx, px[s + 1 < PPL ? s + 1 : s]), __fsub_rn(qy, py[s + 1 < PPL ? s + 1 : s]),
                                   __fsub_rn(qz, pz[s + 1 < PPL ? s + 1 : s])) < r2;
            const unsigned b0 = __ballot_sync(0xffffffffu, h0), b1 = __ballot_sync(0xffffffffu, h1);
            const int pos0 = cnt + __popc(b0 & lt);
            const int mid = cnt + __popc(b0);
            const int pos1 = mid + __popc(b1 & lt);
            if (h0 && pos0 < K) row[pos0] = s * 32 + lane;
            if (h1 && pos1 < K) row[pos1] = (s + 1) * 32 + lane;
            cnt = mid + __popc(b1);
            if (cnt >= K) break;                                     // warp-uniform
        }
        __syncwarp();
        int *out = idx + ci * K;
        if (cnt == 0) {
            if (fill == 1) { for (int l = lane; l < K; l += 32) out[l] = l == 0 ? -1 : 0; }
            else if (fill == 0 && lane == 0) out[0] = -1;   // other slots keep the caller's fill (ball_query_gpu.cu:65)
            // fill == 2: pointnet2_batch semantics, the whole row keeps the caller's fill (no -1 marker)
        } else {
            const int first = row[0];
            for (int l = lane; l < K; l += 32) out[l] = l < cnt ? row[l] : first;
        }
        __syncwarp();                                                // the row is rewritten for the next centroid
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
                                     const int *new_xyz_batch_cnt, const float *xyz, int xyz_stride,
                                     const int *xyz_batch_cnt, int uniform_n, int uniform_m, int *idx, int fill,
                                     void *workspace, size_t workspace_bytes, void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0 && nsample >= 1, "ball_query: bad size B=%d M=%d nsample=%d", B, M, nsample);
    if (M == 0) return MGAR_OK;
    const bool uniform = uniform_n > 0;
    MGAR_REQUIRE(new_xyz && xyz && idx, "ball_query: null pointer");
    MGAR_REQUIRE(uniform || (new_xyz_batch_cnt && xyz_batch_cnt), "ball_query: batch counts missing");
    MGAR_REQUIRE(!uniform || (uniform_m >= 1 && (long long)B * uniform_m == M), "ball_query: uniform layout needs B*uniform_m == M");
    MGAR_REQUIRE(xyz_stride >= 3, "ball_query: xyz_stride must be >= 3");
    cudaStream_t st = (cudaStream_t)stream;
    if (uniform && uniform_n <= 512 && ((size_t)nsample * kBqRegWarps + (size_t)uniform_m * 3) * 4 <= 48 * 1024) {
        const size_t sm = ((size_t)nsample * kBqRegWarps + (size_t)uniform_m * 3) * 4;
#define MGAR_BQ_REG(p)                                                                                            \
    if (uniform_n <= 32 * p) {                                                                                    \
        ball_query_reg_kernel<p><<<B, kBqRegWarps * 32, sm, st>>>(radius, nsample, new_xyz, xyz, xyz_stride,       \
                                                                   uniform_n, uniform_m, idx, fill);               \
        count_launch();                                                                                           \
        return check_launch("ball_query(reg)");                                                                   \
    }
        MGAR_BQ_REG(2) MGAR_BQ_REG(4) MGAR_BQ_REG(8) MGAR_BQ_REG(16)
#undef MGAR_BQ_REG
    }
    const size_t smem = (size_t)kBqChunk * 16 + (size_t)kBqTile * (nsample + 1) * sizeof(int);
    MGAR_REQUIRE(smem <= 200 * 1024, "ball_query: nsample=%d too large", nsample);
    PrefixWs ws;
    int grid;
    if (uniform) {
        grid = B * div_up(uniform_m, kBqTile);
    } else {
        if (int rc = ws.bind(workspace, workspace_bytes, B, "ball_query")) return rc;
        if (int rc = stack_prefix(new_xyz_batch_cnt, xyz_batch_cnt, B, kBqTile, ws.a, ws.b, ws.t, st)) return rc;
        grid = div_up(M, kBqTile) + B;  // upper bound of sum_b ceil(M_b / tile)
    }
    MGAR_CUDA(cudaFuncSetAttribute(ball_query_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ball_query_kernel<<<grid, kBqTile, smem, st>>>(B, radius, nsample, new_xyz, xyz, xyz_stride, ws.a, ws.b, ws.t,
                                                   uniform ? 1 : 0, uniform_n, uniform_m, idx, fill);
    count_launch();
    return check_launch("ball_query");
}

extern "C" int mgar_group_points_stack(int B, int M, int C, int nsample, const float *features,
                                       const int *features_batch_cnt, const int *idx, const int *idx_batch_cnt,
                                       float *out, void *workspace, size_t workspace_bytes, void *stream) {
    MGAR_REQUIRE(B >= 1 && M >= 0 && C >= 0 && nsample >= 0, "group_points: bad size");
    const long long total = (long long)M * C * nsample;
    if (total == 0) return MGAR_OK;
    MGAR_REQUIRE(features && features_batch_cnt && idx && idx_batch_cnt && out, "group_points: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (int rc = ws.bind(workspace, workspace_bytes, B, "group_points")) return rc;
    if (int rc = stack_prefix(idx_batch_cnt, features_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    group_points_kernel<<<div_up(total, 256), 256, 0, st>>>(B, total, C, nsample, features, ws.b, ws.a, idx, out);
    count_launch();
    return check_launch("group_points");
}

extern "C" int mgar_group_points_grad_stack(int B, int M, int C, int N, int nsample, const float *grad_out,
                                            const int *idx, const int *idx_batch_cnt,
                                            const int *features_batch_cnt, float *grad_features, void *workspace,
                                            size_t workspace_bytes, void *stream) {
    (void)N;
    MGAR_REQUIRE(B >= 1 && M >= 0 && C >= 0 && nsample >= 0, "group_points_grad: bad size");
    const long long total = (long long)M * C * nsample;
    if (total == 0) return MGAR_OK;
    MGAR_REQUIRE(grad_out && features_batch_cnt && idx && idx_batch_cnt && grad_features, "group_points_grad: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    PrefixWs ws;
    if (int rc = ws.bind(workspace, workspace_bytes, B, "group_points_grad")) return rc;
    if (int rc = stack_prefix(idx_batch_cnt, features_batch_cnt, B, 1, ws.a, ws.b, nullptr, st)) return rc;
    group_points_grad_kernel<<<div_up(total, 256), 256, 0, st>>>(B, total, C, nsample, grad_out, ws.b, ws.a, idx, grad_features);
    count_launch();
    return check_launch("group_points_grad");
}
