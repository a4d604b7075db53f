// Voxel query: the neighbours of a query point are looked up through a dense voxel -> point index
// map instead of a scan over the cloud.
//
// Reference: pointnet2_stack/src/voxel_query_gpu.cu:10-89 - ONE THREAD PER QUERY walks the (2*zr+1)(2*yr+1)(2*xr+1)
// neighbouring voxels serially; every step is two dependent global loads (map entry -> point xyz), up to 729 in a
// row for the usual [4,4,4] range, with nothing to overlap them but other threads.
// Here: one warp per query.  The neighbourhood is linearised in the reference's (dz, dy, dx) order and the lanes
// take 32 consecutive cells at a time, so the two loads of 32 cells are in flight together; hits are appended in
// order with ballot/popc, the walk stops at the first chunk that fills the row.  Same rows bit for bit: the first
// nsample hits in walk order, padded with the first hit, idx[0] = -1 for a query without neighbours (the other
// slots of such a row are left as the caller zero-filled them, voxel_query_utils.py:33).
#include "mgar_common.cuh"

namespace mgar {

constexpr int kVqWarps = 8;

__global__ void __launch_bounds__(kVqWarps * 32)
voxel_query_stack_kernel(int M, int R1, int R2, int R3, int nsample, float radius, int zr, int yr, int xr,
                         const float *__restrict__ new_xyz, const float *__restrict__ xyz,
                         const int *__restrict__ new_coords, const int *__restrict__ point_indices, int *__restrict__ idx) {
    const int lane = threadIdx.x & 31;
    const long long q = (long long)blockIdx.x * kVqWarps + (threadIdx.x >> 5);
    if (q >= M) return;
    const float r2 = __fmul_rn(radius, radius);
    const float qx = new_xyz[q * 3], qy = new_xyz[q * 3 + 1], qz = new_xyz[q * 3 + 2];
    const int4 c = *reinterpret_cast<const int4 *>(new_coords + q * 4);     // (batch, z, y, x): 16-byte rows
    const int Yn = 2 * yr + 1, Xn = 2 * xr + 1, YX = Yn * Xn;
    const int cells = (2 * zr + 1) * YX;
    const size_t base = (size_t)c.x * R1 * R2 * R3;
    int *row = idx + q * nsample;
    const unsigned lt = (1u << lane) - 1u;
    int cnt = 0, first = -1;
    for (int j0 = 0; j0 < cells && cnt < nsample; j0 += 32) {
        const int j = j0 + lane;
        int nb = -1;
        if (j < cells) {
            const int dz = j / YX, rem = j - dz * YX, dy = rem / Xn, dx = rem - dy * Xn;
            const int z = c.y + dz - zr, y = c.z + dy - yr, x = c.w + dx - xr;
            if (z >= 0 && z < R1 && y >= 0 && y < R2 && x >= 0 && x < R3)
                nb = __ldg(point_indices + base + ((size_t)z * R2 + y) * R3 + x);
        }
        bool hit = false;
        if (nb >= 0) {
            const float px = __ldg(xyz + (size_t)nb * 3), py = __ldg(xyz + (size_t)nb * 3 + 1), pz = __ldg(xyz + (size_t)nb * 3 + 2);
            const float d2 = sqdist(__fsub_rn(px, qx), __fsub_rn(py, qy), __fsub_rn(pz, qz));
            hit = !(d2 > r2);                               // the reference skips on '>' (a NaN distance is kept)
        }
        const unsigned ballot = __ballot_sync(0xffffffffu, hit);
        if (ballot) {
            if (first < 0) first = __shfl_sync(0xffffffffu, nb, __ffs(ballot) - 1);
            const int pos = cnt + __popc(ballot & lt);
            if (hit && pos < nsample) row[pos] = nb;
            cnt += __popc(ballot);
        }
    }
    if (cnt == 0) {
        if (lane == 0) row[0] = -1;
    } else {
        for (int l = min(cnt, nsample) + lane; l < nsample; l += 32) row[l] = first;
    }
}

}  // namespace mgar

using namespace mgar;

extern "C" int mgar_voxel_query_stack(int M, int R1, int R2, int R3, int nsample, float radius, int z_range, int y_range,
                                      int x_range, const float *new_xyz, const float *xyz, const int *new_coords,
                                      const int *point_indices, int *idx, void *stream) {
    MGAR_REQUIRE(M >= 0 && R1 >= 0 && R2 >= 0 && R3 >= 0 && nsample >= 1, "voxel_query: bad size (M=%d nsample=%d)", M, nsample);
    MGAR_REQUIRE(z_range >= 0 && y_range >= 0 && x_range >= 0 && z_range < 128 && y_range < 128 && x_range < 128,
                 "voxel_query: range (%d,%d,%d) outside 0..127", z_range, y_range, x_range);
    if (M == 0) return MGAR_OK;
    MGAR_REQUIRE(new_xyz && xyz && new_coords && point_indices && idx, "voxel_query: null pointer");
    MGAR_REQUIRE((reinterpret_cast<uintptr_t>(new_coords) & 15) == 0, "voxel_query: new_coords must be 16-byte aligned");
    voxel_query_stack_kernel<<<div_up(M, kVqWarps), kVqWarps * 32, 0, (cudaStream_t)stream>>>(
        M, R1, R2, R3, nsample, radius, z_range, y_range, x_range, new_xyz, xyz, new_coords, point_indices, idx);
    count_launch();
    return check_launch("voxel_query");
}
