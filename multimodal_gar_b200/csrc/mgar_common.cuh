// Shared device/host helpers for libmgar_lidar.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mgar_lidar.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libmgar_lidar is written for sm_100a (B200) only"
#endif

namespace mgar {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// ---- error plumbing ------------------------------------------------------------------------
void set_error(const char *fmt, ...);
void count_launch(int n = 1);
int check_launch(const char *what);  // cudaGetLastError -> MGAR_ECUDA + message

#define MGAR_REQUIRE(cond, ...)          \
    do {                                 \
        if (!(cond)) {                   \
            mgar::set_error(__VA_ARGS__); \
            return MGAR_EINVAL;          \
        }                                \
    } while (0)

#define MGAR_CUDA(call)                                                              \
    do {                                                                             \
        cudaError_t e_ = (call);                                                     \
        if (e_ != cudaSuccess) {                                                     \
            mgar::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                            __FILE__, __LINE__);                                     \
            return MGAR_ECUDA;                                                       \
        }                                                                            \
    } while (0)

static inline int div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- in-box arithmetic -----------------------------------------------------------------------
// Pre-processed box: everything that depends on the box alone is computed once per CTA
// instead of once per (point, box) pair as the reference does (roiaware_pool3d_kernel.cu:16-36).
struct BoxP {
    float cx, cy, cz;
    float cosa, sina;  // cosf(-heading), sinf(-heading)
    float hz;          // largest float <= (double)dz/2.0      : in-z  <=> !(|z-cz| > hz)
    float tx, ty;      // smallest float >= dx/2.0 + MARGIN (fp64): in-xy <=> |l| < t
};

// The reference compares float values against fp64 thresholds.  For a float v and a double T,
//   v > T  <=>  v > round_down_f32(T)        v < T  <=>  v < round_up_f32(T)
// so the fp64 comparisons are folded into directed roundings of the per-box thresholds and the
// per-pair work stays in fp32 - bit-identical decisions, no DSETP on the hot loop.
template <bool kCpuSemantics>
__device__ __forceinline__ BoxP make_boxp(const float *b7, const float *cs /* (2) or nullptr */) {
    BoxP p;
    p.cx = b7[0]; p.cy = b7[1]; p.cz = b7[2];
    const float dx = b7[3], dy = b7[4], dz = b7[5], rz = b7[6];
    if (kCpuSemantics) {  // host glibc trig handed in, margin 1e-2 (roiaware_pool3d.cpp:128-140)
        p.cosa = cs[0]; p.sina = cs[1];
    } else {              // libdevice trig, margin 1e-5 (roiaware_pool3d_kernel.cu:16-36)
        p.cosa = cosf(-rz); p.sina = sinf(-rz);
    }
    const double margin = (double)(kCpuSemantics ? 1e-2f : 1e-5f);
    p.hz = __double2float_rd((double)dz / 2.0);
    p.tx = __double2float_ru((double)dx / 2.0 + margin);
    p.ty = __double2float_ru((double)dy / 2.0 + margin);
    return p;
}

// check_pt_in_box3d.  GPU semantics: the contraction nvcc emits for the reference kernels
// (SASS-checked): lx = fma(sx,cos, sy*(-sin)), ly = fma(sy,cos, sx*sin).  CPU semantics: the
// x86-64 host build has no FMA: separate roundings.
template <bool kCpuSemantics>
__device__ __forceinline__ bool in_box(float x, float y, float z, const BoxP &b) {
    if (fabsf(__fsub_rn(z, b.cz)) > b.hz) return false;
    const float sx = __fsub_rn(x, b.cx), sy = __fsub_rn(y, b.cy);
    float lx, ly;
    if (kCpuSemantics) {
        lx = __fadd_rn(__fmul_rn(sx, b.cosa), __fmul_rn(sy, -b.sina));
        ly = __fadd_rn(__fmul_rn(sx, b.sina), __fmul_rn(sy, b.cosa));
    } else {
        lx = __fmaf_rn(sx, b.cosa, __fmul_rn(sy, -b.sina));
        ly = __fmaf_rn(sy, b.cosa, __fmul_rn(sx, b.sina));
    }
    return (fabsf(lx) < b.tx) & (fabsf(ly) < b.ty);
}

// squared distance exactly as the reference kernels compute it once compiled (sampling_gpu.cu:64,
// ball_query_gpu.cu:52; SASS of every instantiation: FMUL on dy, FFMA dx, FFMA dz).
__device__ __forceinline__ float sqdist(float dx, float dy, float dz) {
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

}  // namespace mgar
