// Prefix arrays of ragged (stack-layout) batch counts, shared by the stack ops.
#pragma once
#include "mgar_common.cuh"

namespace mgar {

// pre_a[0..B], pre_b[0..B]: exclusive prefixes of cnt_a / cnt_b; pre_tile[0..B]: exclusive prefix of
// ceil(cnt_a / tile).  Any output may be null.  One small launch on `st`.
int stack_prefix(const int *cnt_a, const int *cnt_b, int B, int tile, int *pre_a, int *pre_b, int *pre_tile,
                 cudaStream_t st);

// Stream-ordered scratch for the three prefix arrays (freed when the object goes out of scope,
// after the consumers have been enqueued on the same stream).
struct PrefixWs {
    int *base = nullptr, *a = nullptr, *b = nullptr, *t = nullptr;
    cudaStream_t st = nullptr;
    int alloc(int B, cudaStream_t s) {
        st = s;
        const size_t n = (size_t)(B + 1);
        cudaError_t e = cudaMallocAsync((void **)&base, 3 * n * sizeof(int), s);
        if (e != cudaSuccess) {
            set_error("stack op: cudaMallocAsync failed: %s", cudaGetErrorString(e));
            base = nullptr;
            return MGAR_ENOMEM;
        }
        a = base; b = base + n; t = base + 2 * n;
        return MGAR_OK;
    }
    ~PrefixWs() { if (base) cudaFreeAsync(base, st); }
};

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

}  // namespace mgar
