// Argument blocks shared by the tensor-core SA-MLP kernels (sa_mlp_tc.cu: 3xTF32, sa_mlp_tc16.cu: 3xFP16).
#pragma once
#include "mgar_common.cuh"

namespace mgar {

constexpr int kTcC1 = 64, kTcC2 = 128, kTcNS = 32;
constexpr int kTcTilePos = 128;                 // positions per tile = 4 centroids x 32 samples

struct TcLayer1 {          // layer-1 parameters travel as kernel arguments (constant bank operands)
    float w[8][kTcC1];     // [c_in][c_out], BatchNorm scale folded in: w * gamma / sqrt(var + eps)
    float sh[kTcC1];       // beta - mean * scale
};
// bound of |layer-1 pre-activation| used by the fp16 kernel to scale a centroid into the fp16 range:
// |z1[k]| <= sum_c wmax[c] * |x[c]| + shmax
struct TcBound { float wmax[8]; float shmax; };

struct TcArgs {
    int B, M, C_in;
    const float *xyz; int xyz_stride;
    const float *new_xyz;
    const float *features; int feat_stride;
    const int *pre_xyz, *pre_new; int uniform, n_u, m_u;
    const int *idx;
    const unsigned char *empty_mask;
    const float *w2t /* [64][128] */, *sc2, *sh2;
    float *out; int ld_out;              // (M, 128) SA features, may be null
    float *gmax; int group_size;         // (ceil(M / group_size), 128) max over each group of centroids, may be null
    const int *group_zero;               // per group: non-zero -> the group's row stays zero (empty person)
#ifdef MGAR_TC_TRACE
    long long *trace;                    // timeline stamps of CTA 0, tiles 100..131 (debug build only)
#endif
    int packed4;                         // xyz/features are columns 0..2 / 3 of one 16B-aligned (N,4) array
    // training-statistics variant (sa_mlp_tc16_launch_train): sc2 holds +-1 = sign(gamma2), sh2 zeros, no BatchNorm /
    // ReLU in the epilogue; per (centroid, channel) the extreme of sign * z2 over the samples and the first sample
    // attaining it, per channel sum z2 and sum z2^2 (fp64 atomics)
    float *tr_zext; int *tr_arg; double *tr_sum, *tr_sq;
    // training variant only (null elsewhere): layer 1 on the DEVICE - 8 x 64 folded weights [c_in][c_out] | 64 shifts |
    // wmax[8], shmax (the TcLayer1 / TcBound contents, written by f_bn1_kernel) - so that the statistics pass needs no
    // device-to-host copy of BatchNorm 1's batch scale / shift (and is capturable into a CUDA graph)
    const float *l1_dev;
    // empty-person skip (sa_mlp_tc16 only; null = off): live_groups[0 .. *n_live) = indices of the groups (persons)
    // whose group_zero flag is 0, ascending.  The kernel then walks only n_live * group_size centroids ("virtual"
    // centroid v -> real centroid live_groups[v / group_size] * group_size + v % group_size).  Needs out == null,
    // gmax != null (zero-filled: skipped persons keep their zero row), the uniform layout with m_u == group_size,
    // group_size % 4 == 0 (a tile of 4 centroids never straddles two persons).
    const int *live_groups, *n_live;
    // distinct-centroid walk (sa_mlp_tc16 only; null = off): live_tiles[0 .. *n_live) = first REAL centroid of every
    // tile of four centroids worth computing - for each person whose group_zero flag is 0, the tiles covering its
    // first live_cnt centroids (farthest point sampling returns index 0 once every distinct point has been taken,
    // so the centroids live_cnt .. group_size-1 repeat centroid 0 and cannot change the person's max).  Same
    // preconditions as live_groups; m_shift = log2(group_size) or -1.
    const int *live_tiles; int m_shift;
    // distinct-sample walk (sa_mlp_tc16 only, ns = 32, out == null; null = off): live_units[U t + u], t < *n_live, u < U =
    // (centroid << log2(32 / s) | unit) [| 1 << 31: empty ball]: tile t is U = 128 / s UNITS of s sample slots (slots
    // s o .. s o + s - 1 of the centroid's ball row; s = 4, or 8 - "octets" - in the first version) of ONE person, instead
    // of four whole centroids.  A ball row repeats its hits once it runs out
    // of distinct ones (copies of a wrapped cloud, then the first hit as padding); the list holds only the octets that
    // start with a distinct sample.  A TMEM lane quarter (32 positions) then mixes up to four centroids: the epilogue's
    // max over the quarter, the folded BatchNorm + ReLU (monotone) and the running person max still commute with the
    // reference's max over a centroid's samples followed by the max over the person's centroids; the power-of-two
    // operand scale is per quarter either way.
    const int *live_units;
    // samples per centroid: 32 (a TMEM lane quarter = one centroid), or - sa_mlp_tc16 with out == null only - 16 / 64.
    // The kernel always walks BLOCKS of 32 consecutive positions of the flat (M * ns) sample array: a block is two
    // centroids (ns = 16) or half a centroid (ns = 64), and M, m_u, group_size, live_tiles, m_shift count blocks.  The max
    // over a block, the folded BatchNorm + ReLU (monotone) and the max over a person's blocks commute with the
    // reference's max over a centroid's samples followed by the max over the person's centroids.
    int ns;
};

// sa_mlp_tc16.cu: the 3xFP16 kernel (default); sa_mlp_tc.cu keeps the 3xTF32 kernel (MGAR_TC_MODE=tf32)
int sa_mlp_tc16_launch(const TcLayer1 &L1, const TcArgs &A, int C0, int grid, cudaStream_t st);
// live[0 .. *n_live) = ascending indices g < G with group_zero[g] == 0 (one small launch)
int tc_compact_live_groups(const int *group_zero, int G, int *live, int *n_live, cudaStream_t st);
// tiles[0 .. *n_tiles) = g * group_size + 4 q for every g < G with group_zero[g] == 0 and q < ceil(blocks / 4), blocks =
// ceil(live_cnt[g] * ns / 32) position blocks of 32 (group_size counts blocks), ascending (one small launch)
int tc_build_live_tiles(const int *group_zero, const int *live_cnt, int G, int group_size, int ns, int *tiles, int *n_tiles,
                        cudaStream_t st);
// units[16 t + u] as TcArgs::live_units describes, *n_tiles = number of tiles (zeroed here; one small launch).  point_cnt[g] =
// number of DISTINCT points of group g's cloud (slot k of the cloud holds point k mod point_cnt[g]).
int tc_units_per_row();        // list entries a centroid can take (32 / slots per entry)
int tc_units_per_tile();       // entries per 128-position tile
int tc_build_live_units(const int *group_zero, const int *live_cnt, const int *point_cnt, const int *idx, int G, int group_size,
                        int n_u, int *units, int *n_tiles, cudaStream_t st);
int sa_mlp_tc16_launch_train(const TcLayer1 &L1, const TcArgs &A, int C0, int grid, cudaStream_t st);   // C0 = 4, 5

}  // namespace mgar
