// Launchers of the image-primitive kernels (reference: src/testbed_image.cu).
#pragma once
#include "ngp_common.cuh"
namespace ngp {
enum ImageRandomMode : int { IMG_RANDOM_RANDOM = 0, IMG_RANDOM_HALTON = 1, IMG_RANDOM_SOBOL = 2, IMG_RANDOM_STRATIFIED = 3 };   // ERandomMode (common.h:124-130)
// base_idx = batch_size * training_step and seed feed the Halton / Sobol sequences; rng feeds Random / Stratified
void image_generate_batch(cudaStream_t s, uint32_t n, Pcg32 rng, int random_mode, size_t base_idx, uint32_t seed, const float* image_rgba, int W, int H, bool snap,
                          bool linear_colors, float* pos /*[n][2]*/, float* target /*[n][3]*/);
void image_eval_coords(cudaStream_t s, uint32_t n, uint32_t offset, const float* image_rgba, int W, int H, bool linear_colors, float* pos, float* target);
void image_from_rgba8(cudaStream_t s, uint64_t n_pixels, const uint8_t* rgba8, float* out_rgba);
void image_squared_error(cudaStream_t s, uint32_t n, const float* target, const __half* pred, float* se, bool quantize);
}
