// Entry points of the specialised training path (sa_mlp_train_fast.cu), called from sa_mlp_train.cu.
#pragma once
#include "mgar_common.cuh"

namespace mgar {
namespace trfast {

// shape test: two layers [3+C_in, 64, 128], nsample 32, max-pool, C_in 1 or 2 (and MGAR_SA_TRAIN_KERNELS != generic)
bool applicable(int M, int nsample, int n_layers, const int *channels, int pool_avg);
size_t workspace_bytes(int M, int C0);

int forward(int B, int M, int C_in, const float *xyz, const int *xyz_batch_cnt, const float *new_xyz,
            const int *new_xyz_batch_cnt, int uniform_n, int uniform_m, const float *features, const int *idx,
            const unsigned char *empty_mask, const float *weights, const float *gamma_beta, float eps, float *out,
            int ld_out, float *batch_stats, void *workspace, size_t ws_bytes, cudaStream_t st);

int backward(int B, int M, int C_in, const int *xyz_batch_cnt, const int *new_xyz_batch_cnt, int uniform_n,
             int uniform_m, const int *idx, const unsigned char *empty_mask, const float *weights,
             const float *gamma_beta, const float *grad_out, int ld_grad, void *workspace, size_t ws_bytes,
             float *grad_weights, float *grad_gamma_beta, float *grad_features, cudaStream_t st);

}  // namespace trfast
}  // namespace mgar
