// Internal C++ interface between the host objects (testbed.cu, capi.cu) and the kernel launchers.
// Nothing here is exported; the public boundary is include/ngp_b200.h.
#pragma once
#include "ngp_common.cuh"

namespace ngp {

// ------------------------------------------------------------------------------------------------
// Multiresolution hash grid (replaces tcnn::GridEncoding<__half>; reference call sites
// nerf_network.h:93-95,143-148,201-207,324-333,345-350, testbed.cu:3961-4005,4101)
// ------------------------------------------------------------------------------------------------
constexpr uint32_t GRID_MAX_LEVELS = 32;
constexpr uint32_t GRID_PRIME_1 = 2654435761u, GRID_PRIME_2 = 805459861u;
constexpr uint32_t GRID_PARAM_ALIGN = 8;
constexpr float GRID_INIT_RANGE = 1e-4f;

enum ActLayout : int { LAYOUT_ROWMAJOR = 0, LAYOUT_BLK8 = 1 };

struct GridLayout {
    uint32_t n_dims, n_levels, n_features, log2_hashmap_size, base_resolution;
    float per_level_scale;
    uint32_t padded_width;                       // C: L*F rounded up to 16 (FullyFusedMLP alignment)
    uint32_t hashed_mask;                        // bit l set: level l uses the prime-xor hash
    uint32_t offsets[GRID_MAX_LEVELS + 1];       // entry offsets (x F for element offsets)
    uint32_t sizes[GRID_MAX_LEVELS];
    uint32_t resolutions[GRID_MAX_LEVELS];
    float scales[GRID_MAX_LEVELS];
    uint32_t dense_stride[GRID_MAX_LEVELS][3];   // uint32-wrapping strides of the dense index (0: dim not used)
    uint64_t n_params;
};

void grid_make_layout(GridLayout& g, uint32_t n_dims, uint32_t n_levels, uint32_t n_features, uint32_t log2_hashmap_size,
                      uint32_t base_resolution, float per_level_scale);
// y: [n_rows_padded x C] fp16 in `y_layout`; rows [0,n) are written (n_levels*F.. C zero-filled)
// n_dyn (optional, device memory): the kernels process min(n, *n_dyn) rows; n sizes the launch.  Used by the inference tracer, whose row
// count per iteration is produced on the device by the ray-compaction kernel
void grid_forward(cudaStream_t s, const GridLayout& g, const __half* table, const float* x, uint32_t x_stride, uint32_t n,
                  __half* y, int y_layout, const uint32_t* n_dyn = nullptr, const float* max_level = nullptr);
// max_level (optional, device memory, one float per row; GridEncoding::set_max_level_gpu of tiny-cuda-nn, recalled - the source is not in
// the mount): row i encodes only the levels l with l < max_level[i] * n_levels + 1e-3 (the others give zeros) and, backward, scatters
// gradients only to the levels with l <= max_level[i] * n_levels + 1e-3.  1.0 and above = every level.
// grad += scatter(w * dy); grad is table-shaped fp16 (caller zeroes it for EGradientMode::Overwrite)
void grid_backward(cudaStream_t s, const GridLayout& g, const float* x, uint32_t x_stride, uint32_t n, const __half* dy,
                   int dy_layout, __half* grad, const float* max_level = nullptr);
void grid_backward_chunks(cudaStream_t s, const GridLayout& g, const float* x, uint32_t x_stride, uint32_t n, const __half* dy,
                          int dy_layout, __half* grad, uint32_t chunk0, uint32_t n_chunks, const float* max_level = nullptr);
// dL/dx [n][n_dims] fp32
void grid_backward_input(cudaStream_t s, const GridLayout& g, const __half* table, const float* x, uint32_t x_stride,
                         uint32_t n, const __half* dy, int dy_layout, float* dx, uint32_t dx_stride = 0, const float* max_level = nullptr);   // dx_stride 0: n_dims
// debug / parity: level-local corner indices [n][2^D] and weights for one level
void grid_corners(cudaStream_t s, const GridLayout& g, const float* x, uint32_t x_stride, uint32_t n, uint32_t level,
                  uint32_t* idx, float* w);
// params[i] = U(-1e-4, 1e-4) from float #i of rng (counter based)
void grid_init(cudaStream_t s, const GridLayout& g, Pcg32 rng, float* params_fp32);

// ------------------------------------------------------------------------------------------------
// Fully fused 64-wide MLPs on tcgen05/TMEM (replaces tcnn::FullyFusedMLP<__half,64> and the NerfNetwork glue;
// reference: nerf_network.h:76-578).  Weights: fp16 row-major [out x in] per layer, layers concatenated.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t MLP_WIDTH = 64;
constexpr uint32_t MLP_OUT_PAD = 16;

struct NerfNetDims {
    uint32_t pos_width;        // padded encoder width: 16 or 32
    uint32_t density_hidden;   // 1 or 2
    uint32_t rgb_hidden;       // 1, 2 or 3
};
inline uint32_t mlp_n_params(uint32_t in_w, uint32_t n_hidden) { return MLP_WIDTH * in_w + (n_hidden - 1) * MLP_WIDTH * MLP_WIDTH + MLP_OUT_PAD * MLP_WIDTH; }

// Inference of the whole NeRF network on n (multiple of 128) samples.
//   enc: [n x pos_width] blk8 fp16 ; coords: [n][7] fp32 (dir at +4) ; out: [n][4] fp16 (r,g,b,sigma raw)
void nerf_mlp_forward(cudaStream_t s, const NerfNetDims& d, const __half* w_density, const __half* w_rgb, const __half* enc,
                      const float* coords, uint32_t n, __half* out, const uint32_t* n_dyn = nullptr);
// density net only: out [n] fp16 (channel 0)
void nerf_mlp_density(cudaStream_t s, const NerfNetDims& d, const __half* w_density, const __half* enc, uint32_t n, __half* out);
// Forward recompute + backward on n samples.  dL_dout: [n][4] fp16.  Outputs: dL_denc [n x pos_width] blk8 fp16,
// wgrad_f32: [n_density_params + n_rgb_params] fp32, ACCUMULATED with atomics (caller zeroes).
void nerf_mlp_backward(cudaStream_t s, const NerfNetDims& d, const __half* w_density, const __half* w_rgb, const __half* enc,
                       const float* coords, uint32_t n, const __half* dL_dout, __half* dL_denc, float* wgrad_f32, float* dL_dcoords = nullptr);
// dL_dcoords (optional): [n][7] fp32 input gradients; the kernel writes the direction part (cols 4..6) = dL/d(SH4 encoding) x dSH4/d(dir)

// Plain MLP (image / SDF mode: NetworkWithInputEncoding, testbed.cu:4110): in_w in {16,32}, n_hidden in {1,2,3}
//   out: [n][out_stride] fp16, first 16 columns written
void mlp_forward(cudaStream_t s, uint32_t in_w, uint32_t n_hidden, const __half* w, const __half* in_blk8, uint32_t n,
                 __half* out, uint32_t out_stride);
void mlp_backward(cudaStream_t s, uint32_t in_w, uint32_t n_hidden, const __half* w, const __half* in_blk8, uint32_t n,
                  const __half* dL_dout, uint32_t dout_stride, __half* dL_din_blk8, float* wgrad_f32);

// ------------------------------------------------------------------------------------------------
// Optimizer (replaces tcnn Ema(ExponentialDecay(Adam)); reference: testbed_nerf.cu:3678, configs/nerf/base.json:5-22)
// ------------------------------------------------------------------------------------------------
struct AdamHyper {
    float learning_rate, beta1, beta2, epsilon, l2_reg, loss_scale;
    uint64_t n_matrix_params;
    int optimize_matrix, optimize_non_matrix;
    float ema_decay;      // 0: no EMA
    uint32_t ema_step;    // 1-based step for debiasing
    int clear_consumed_grads;   // write zeros back over the non-zero hash-table gradients once read: the table is then clean for the
                                // next backward pass, which can skip its 24 MB memset (Trainer; the C ABI entry leaves gradients alone)
};
// grads_f32_matrix (optional): fp32 gradients of the first n_matrix params (MLP wgrad staging); if null, fp16 grads used
void adam_step(cudaStream_t s, const AdamHyper& h, uint64_t n, float* master, __half* params, const __half* grads,
               float* m, float* v, uint32_t* steps, float* ema_f32, __half* ema_half);
void cast_f32_to_half(cudaStream_t s, const float* in, __half* out, uint64_t n);

// Roofline denominators measured on this device (microbench.cu): random 4-byte gathers per second over an L2-sized buffer,
// scattered fp16x2 reduction lanes per second (vec = 4-byte entries per lane: 1, 2 or 4)
double microbench_l2_gather(cudaStream_t s, uint64_t buffer_bytes, uint32_t reps);
double microbench_l2_red(cudaStream_t s, uint64_t buffer_bytes, int vec, uint32_t reps);

}  // namespace ngp
