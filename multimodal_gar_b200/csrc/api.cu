// Error plumbing and bookkeeping shared by every entry point of libmgar_lidar.so.
#include <stdarg.h>

#include <atomic>

#include "mgar_common.cuh"
#include "stack_prefix.cuh"

namespace mgar {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

int check_launch(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: kernel launch failed: %s", what, cudaGetErrorString(e));
        return MGAR_ECUDA;
    }
    return MGAR_OK;
}

}  // namespace mgar

extern "C" {

int mgar_abi_version(void) { return MGAR_ABI_VERSION; }
const char *mgar_last_error(void) { return mgar::g_err; }
unsigned long long mgar_launch_count(void) { return mgar::g_launches.load(); }
size_t mgar_stack_workspace_bytes(int B) { return B > 0 ? mgar::stack_ws_bytes(B) : 0; }

// Pinned HOST staging buffers for the host-buffer front end (pipeline.HostFramePipeline): page-locked, portable across
// devices, optionally write-combined (the CPU only ever WRITES frames into them; the copy engines read them without
// snooping the CPU caches).  Explicit, caller-driven: nothing is allocated behind an operator call.
int mgar_host_alloc(void **ptr, size_t bytes, int write_combined) {
    MGAR_REQUIRE(ptr && bytes > 0, "host_alloc: bad argument");
    MGAR_CUDA(cudaHostAlloc(ptr, bytes, cudaHostAllocPortable | (write_combined ? cudaHostAllocWriteCombined : 0)));
    return MGAR_OK;
}
int mgar_host_free(void *ptr) {
    if (ptr) MGAR_CUDA(cudaFreeHost(ptr));
    return MGAR_OK;
}

}  // extern "C"
