// Prefix arrays of ragged (stack-layout) batch counts, shared by the stack ops.
#pragma once
#include "mgar_common.cuh"

namespace mgar {

// pre_a[0..B], pre_b[0..B]: exclusive prefixes of cnt_a / cnt_b; pre_tile[0..B]: exclusive prefix of
// ceil(cnt_a / tile).  Any output may be null.  One small launch on `st`.
int stack_prefix(const int *cnt_a, const int *cnt_b, int B, int tile, int *pre_a, int *pre_b, int *pre_tile,
                 cudaStream_t st);

// The three prefix arrays live in CALLER-PROVIDED scratch (the binding passes workspace pointers; the
// library allocates nothing): mgar_stack_workspace_bytes(B) bytes at `base`.  (Round 1 took them from
// cudaMallocAsync per call: the pool hands its memory back at every synchronisation, so each ragged call paid a fresh
// ~4 ms allocation - 26 of the 53 ms of a training step.)
constexpr size_t stack_ws_bytes(int B) { return (3 * ((size_t)B + 1) * sizeof(int) + 15) & ~(size_t)15; }
struct PrefixWs {
    int *a = nullptr, *b = nullptr, *t = nullptr;
    int bind(void *base, size_t avail, int B, const char *who) {
        if (!base || avail < stack_ws_bytes(B)) {
            set_error("%s: ragged (stack-layout) batches need a workspace of mgar_stack_workspace_bytes(B=%d) = %zu bytes, got %zu",
                      who, B, stack_ws_bytes(B), base ? avail : (size_t)0);
            return MGAR_ENOMEM;
        }
        const size_t n = (size_t)(B + 1);
        a = static_cast<int *>(base); b = a + n; t = a + 2 * n;
        return MGAR_OK;
    }
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
