// Launchers of the image-primitive kernels (reference: src/testbed_image.cu).
#pragma once
#include "ngp_common.cuh"
namespace ngp {
void image_generate_batch(cudaStream_t s, uint32_t n, Pcg32 rng, bool stratify, const float* image_rgba, int W, int H, bool snap, bool linear_colors,
                          float* pos /*[n][2]*/, float* target /*[n][3]*/);
void image_eval_coords(cudaStream_t s, uint32_t n, uint32_t offset, const float* image_rgba, int W, int H, bool linear_colors, float* pos, float* target);
void image_squared_error(cudaStream_t s, uint32_t n, const float* target, const __half* pred, float* se, bool quantize);
}
