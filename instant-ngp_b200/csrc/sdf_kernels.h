// Launchers of the SDF-primitive training kernels (reference: src/testbed_sdf.cu:1289-1311 train_sdf, and the tcnn pieces it
// calls: shuffle<T> + permute, create_loss "MAPE" / "L2" / "RelativeL2" / "L1" / "SMAPE").
#pragma once
#include "ngp_common.cuh"
namespace ngp {

// tcnn::create_loss otypes that Trainer::training_step applies element-wise (image and SDF modes)
enum TcnnLoss : int { TLOSS_L2 = 0, TLOSS_RELATIVE_L2 = 1, TLOSS_L1 = 2, TLOSS_MAPE = 3, TLOSS_SMAPE = 4 };

// tcnn shuffle<T>(n_elements, stride=1, seed, in, out): out[i] = in[permute(i, n_elements, seed)], positions (3 floats) and
// distances in one pass; only the first n_out records of the permutation are produced (train_sdf consumes min(size, batch))
void sdf_shuffle(cudaStream_t s, uint32_t n_elements, uint32_t n_out, uint32_t seed, const float* positions, const float* distances,
                 float* positions_shuffled, float* distances_shuffled);
// element-wise tcnn loss: pred fp16 [n][16] (first `dims` columns used), target fp32 [n][dims]; dL fp16 [n][16] (padding columns zero),
// loss_values[n] = sum over dims of the per-element value (already divided by n*dims); loss_values may be null
void tcnn_loss(cudaStream_t s, int loss_type, uint32_t n, uint32_t dims, const __half* pred, const float* target, float loss_scale, __half* dL,
               float* loss_values);
}  // namespace ngp
