// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// Scalar CPU restatement of the instant-ngp hot path (yblin/instant-ngp @ b5ba2a3) used as the
// parity checker for the CUDA kernels in instant-ngp_b200/csrc.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load this library.  The product path never
// links or calls it and has no CPU fallback.
//
// PARITY STATUS: "parity unpinned" for everything that lives in tiny-cuda-nn
// (NVlabs/tiny-cuda-nn @ ef04dfd2713916d218323027fe6f0194f2b7d16d, an EMPTY submodule in the
// reference mount): hash-grid indexing/interpolation, FullyFusedMLP numerics,
// SH basis, Adam/EMA/ExponentialDecay, pcg32, morton3D, fill_rollover.  Those are restated from
// the published algorithm (Müller et al. 2022) and the reference's call sites; every convention
// is a named constant below so it can be re-pinned if tcnn sources appear.
// The in-mount parts (sampler, occupancy grid, loss/compositor, tracer, image step) follow
// /root/reference/src/testbed_nerf.cu, testbed_image.cu, include/.../{nerf_network.h,
// common_device.cuh,random_val.cuh,bounding_box.cuh} line by line (citations at each function),
// and are cross-checked against the reference's own __host__ __device__ helpers compiled into
// oracle/_ref (see oracle/Makefile, oracle/ref_shim/).
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

// ---------------------------------------------------------------------------------------------
// Conventions recalled from tiny-cuda-nn (unpinned; see header).
// ---------------------------------------------------------------------------------------------
#define REF_GRID_PRIME_1 2654435761u   // coherent_prime_hash primes {1, P1, P2, ...}
#define REF_GRID_PRIME_2 805459861u
#define REF_GRID_PRIME_3 3674653429u
#define REF_GRID_PARAM_ALIGN 8u        // params_in_level rounded up to a multiple of 8
#define REF_GRID_INIT_RANGE 1e-4f      // U(-1e-4, 1e-4); same constant in-mount: takikawa_encoding.cuh:429
#define REF_BATCH_GRANULARITY 256u     // tcnn::batch_size_granularity (testbed_image.cpp:300,480 uses 256)
#define REF_MLP_OUT_PAD 16u            // FullyFusedMLP pads n_output_dims to 16

typedef struct RefGridCfg {
    uint32_t n_dims;              // 2 or 3
    uint32_t n_levels;            // L
    uint32_t n_features;          // F per level
    uint32_t log2_hashmap_size;   // log2 T
    uint32_t base_resolution;     // N_min
    float    per_level_scale;     // b
} RefGridCfg;

// -- RNG (pcg32, O'Neill; tcnn::default_rng_t) ---------------------------------------------------
typedef struct RefPcg32 { uint64_t state, inc; } RefPcg32;
void     ref_pcg32_seed(RefPcg32* r, uint64_t initstate, uint64_t initseq);
uint32_t ref_pcg32_next_uint(RefPcg32* r);
float    ref_pcg32_next_float(RefPcg32* r);
void     ref_pcg32_advance(RefPcg32* r, int64_t delta);
// out[i] = i-th float of stream pcg32(seed) advanced by `skip`
void     ref_pcg32_fill_float(uint64_t seed, uint64_t skip, float* out, size_t n);
void     ref_pcg32_fill_uint(uint64_t seed, uint64_t skip, uint32_t* out, size_t n);

// -- fp16 helpers ----------------------------------------------------------------------------------
void ref_float_to_half(const float* in, uint16_t* out, size_t n);
void ref_half_to_float(const uint16_t* in, float* out, size_t n);

// -- hash grid -------------------------------------------------------------------------------------
// offsets: [L+1] entry offsets (not multiplied by F); returns total #params = offsets[L]*F
uint64_t ref_grid_layout(const RefGridCfg* cfg, uint32_t* offsets, float* scales, uint32_t* resolutions);
// per (point, level): 2^D entry indices (level-local, before adding the offset) and weights
void ref_grid_corners(const RefGridCfg* cfg, const float* x, uint32_t x_stride, uint32_t n, uint32_t level,
                      uint32_t* idx_out, float* w_out);
// per-row level cut-off for the ref_grid_forward / ref_grid_backward calls that follow (GridEncoding::set_max_level_gpu, recalled); NULL = off
void ref_grid_set_max_level(const float* max_level);
// y[n][y_stride] fp16 (columns >= L*F up to y_width are zero-filled)
void ref_grid_forward(const RefGridCfg* cfg, const uint16_t* table, const float* x, uint32_t x_stride, uint32_t n,
                      uint16_t* y, uint32_t y_stride, uint32_t y_width);
// grad (double, table-shaped, accumulated +=) from dy[n][dy_stride] fp16
void ref_grid_backward(const RefGridCfg* cfg, const float* x, uint32_t x_stride, uint32_t n,
                       const uint16_t* dy, uint32_t dy_stride, double* grad);
// dy/dx: dx[n][D] float
void ref_grid_backward_input(const RefGridCfg* cfg, const uint16_t* table, const float* x, uint32_t x_stride,
                             uint32_t n, const uint16_t* dy, uint32_t dy_stride, float* dx);
void ref_grid_init(const RefGridCfg* cfg, uint64_t seed, float* params_fp32);

// -- spherical harmonics degree 4 on 2*d-1 (dir in [0,1]^3) -----------------------------------------
void ref_sh4(const float* dirs, uint32_t stride, uint32_t n, uint16_t* out, uint32_t out_stride);

// -- fully fused MLP (ReLU hidden, linear out, fp16 storage, fp32 accumulate) -----------------------
// weights: fp16, row-major per layer [out x in]: first [W x in_width], (n_hidden-1) x [W x W], last [16 x W]
uint32_t ref_mlp_n_params(uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t out_pad);
// acts (optional): [n_hidden][n][width] fp16 post-activation
void ref_mlp_forward(const uint16_t* weights, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t out_pad,
                     const uint16_t* in, uint32_t in_stride, uint32_t n, uint16_t* out, uint32_t out_stride,
                     uint16_t* acts);
// dL_dout: [n][dout_stride] fp16 (out_pad used); dL_din (optional) [n][din_stride] fp16; wgrad double (+=)
void ref_mlp_backward(const uint16_t* weights, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t out_pad,
                      const uint16_t* in, uint32_t in_stride, uint32_t n, const uint16_t* acts,
                      const uint16_t* dL_dout, uint32_t dout_stride, uint16_t* dL_din, uint32_t din_stride,
                      double* wgrad);
void ref_mlp_init(uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t out_pad, RefPcg32* rng, float* params);

// -- NerfNetwork (nerf_network.h:76-578) -------------------------------------------------------------
typedef struct RefNerfNetCfg {
    RefGridCfg grid;
    uint32_t width;             // 64
    uint32_t density_hidden;    // n_hidden_layers of density net
    uint32_t rgb_hidden;        // n_hidden_layers of rgb net
} RefNerfNetCfg;
// per-sample level cut-offs of the position encoding for the ref_nerf_forward / ref_nerf_backward calls that follow
// (max_level_rand_training: GridEncoding::set_max_level_gpu, testbed_nerf.cu:3996,4004); NULL = off
void ref_nerf_set_max_level(const float* max_level);
uint32_t ref_nerf_pos_width(const RefNerfNetCfg* c);      // padded encoder width (multiple of 16)
uint64_t ref_nerf_n_params(const RefNerfNetCfg* c, uint64_t* off_density, uint64_t* off_rgb, uint64_t* off_grid);
// coords: [n][7] float (pos3, dt, dir3) ; out: [n][4] fp16 (r,g,b,sigma raw)
void ref_nerf_forward(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n, uint16_t* out);
// density only: pos [n][pos_stride]; out [n] fp16 = density-net output channel 0
void ref_nerf_density(const RefNerfNetCfg* c, const uint16_t* params, const float* pos, uint32_t pos_stride,
                      uint32_t n, uint16_t* out);
// full fwd+bwd: dL_dout [n][4] fp16 -> grads (double, params-shaped, +=)
void ref_nerf_backward(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n,
                       const uint16_t* dL_dout, double* grads);
// the same, also returning dL/d(coords) [n][7] fp32 (NerfNetwork::backward's dL_dinput: position cols 0..2, direction cols 4..6)
void ref_nerf_backward_input(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n,
                             const uint16_t* dL_dout, double* grads, float* dL_dcoords);
// d(SH4 output k)/d(direction input axis a), J[16][3] row-major, for one direction in [0,1]^3
void ref_sh4_jacobian(const float* dir3, float* J48);
void ref_nerf_init(const RefNerfNetCfg* c, uint64_t seed, float* params_fp32);

// -- optimizer: Ema(ExponentialDecay(Adam)) (configs/nerf/base.json:5-22) -----------------------------
typedef struct RefAdamCfg {
    float learning_rate, beta1, beta2, epsilon, l2_reg;
    float loss_scale;
    uint64_t n_matrix_params;       // params [0, n_matrix) are MLP weights (L2 reg, always stepped)
    int optimize_matrix, optimize_non_matrix;
} RefAdamCfg;
void ref_adam_step(const RefAdamCfg* c, uint64_t n, float* master, uint16_t* params, const uint16_t* grads,
                   float* m, float* v, uint32_t* steps);
float ref_exp_decay_lr(float base_lr, uint32_t step, uint32_t decay_start, uint32_t decay_interval, float decay_base,
                       uint32_t decay_end);
void ref_ema_step(uint64_t n, float decay, uint32_t step, const uint16_t* params, float* ema_fp32, uint16_t* ema_half);

// ---------------------------------------------------------------------------------------------
// NeRF sampler / occupancy grid / loss (oracle/ngp_ref_nerf.cpp) — declared in ngp_ref_nerf.h
// ---------------------------------------------------------------------------------------------
#ifdef __cplusplus
}
#endif
