// Host-side model objects mirroring the reference's operator surface:
//   HashGridEncoding      <-> tcnn::create_encoding<T>(n_dims, json, alignment)        (nerf_network.h:93-95, testbed.cu:4101)
//   NerfNetwork           <-> ngp::NerfNetwork<T>                                       (nerf_network.h:76-578)
//   EncodedMlp            <-> tcnn::NetworkWithInputEncoding<T>                         (testbed.cu:4110; image / SDF modes)
//   Optimizer             <-> tcnn::create_optimizer(json): Ema(ExponentialDecay(Adam)) (testbed.cu:4008, configs/nerf/base.json:5-22)
//   Trainer               <-> tcnn::Trainer<float,T,T> (owns master/params/inference params/gradients)  (testbed.cu:4129)
#pragma once
#include <functional>
#include "ngp_internal.h"
#include "json_mini.h"
#include <vector>

namespace ngp {

// grow-only device scratch (stands in for tcnn::GPUMemoryArena allocations tied to a stream)
struct DeviceBuffer {
    void* ptr = nullptr; size_t bytes = 0;
    DeviceBuffer() = default;
    DeviceBuffer(const DeviceBuffer&) = delete;
    DeviceBuffer& operator=(const DeviceBuffer&) = delete;
    ~DeviceBuffer() { if (ptr) cudaFree(ptr); }
    template <class T> T* get(size_t count) {
        size_t need = count * sizeof(T);
        if (need > bytes) { if (ptr) NGP_CUDA_CHECK(cudaFree(ptr)); ptr = nullptr; bytes = 0; size_t cap = need + need / 4 + 256; NGP_CUDA_CHECK(cudaMalloc(&ptr, cap)); bytes = cap; }
        return reinterpret_cast<T*>(ptr);
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(ptr); }
};

struct MlpConfig { uint32_t n_neurons = 64, n_hidden_layers = 2, n_output_dims = 16; };
MlpConfig parse_mlp_config(const Json& j, const char* what);
void grid_from_json(GridLayout& g, const Json& enc, uint32_t n_pos_dims);
void mlp_init_host(uint32_t in_w, uint32_t n_hidden, Pcg32& rng, float* params);   // Xavier uniform, memory order

struct NerfNetwork {
    GridLayout grid; NerfNetDims dims;
    uint64_t off_density = 0, off_rgb = 0, off_grid = 0, n_params = 0, n_matrix_params = 0;
    DeviceBuffer enc_buf, denc_buf, wgrad_buf, enc_train_buf;
    // GridEncoding::set_max_level_gpu (testbed_nerf.cu:3996,4004,4130): per-row level cut-off of the position encoding for the calls that
    // follow (one float per sample row; nullptr = all levels).  Used by max_level_rand_training only.
    const float* max_level_gpu = nullptr;

    // config = the reference's network JSON ("encoding", "network", "dir_encoding", "rgb_network")
    explicit NerfNetwork(const Json& config);
    void initialize_params(uint64_t seed, float* master_dev, cudaStream_t s);
    // coords [n][7] fp32; out [n][4] fp16 ; n % 128 == 0
    void inference(cudaStream_t s, const __half* params, const float* coords, uint32_t n, __half* out, const uint32_t* n_dyn = nullptr);
    // pos [n][pos_stride] fp32; out [n] fp16 (raw density)
    void density(cudaStream_t s, const __half* params, const float* pos, uint32_t pos_stride, uint32_t n, __half* out);
    // forward(recompute)+backward; grads: param-shaped fp16, OVERWRITTEN (EGradientMode::Overwrite, testbed_nerf.cu:4078)
    // dL_dcoords (optional): [n][7] fp32, the gradient w.r.t. the input rows (NerfNetwork::backward's dL_dinput, nerf_network.h:282-330): position
    // cols 0..2 through the hash grid, direction cols 4..6 through SH4; col 3 (dt) is not an input of the network and is left alone
    void forward_backward(cudaStream_t s, const __half* params, const float* coords, uint32_t n, const __half* dL_dout, __half* grads, float* dL_dcoords = nullptr);
    // same, with the encoder outputs of the n samples already known (blk8 rows in enc_train_buf, carried over from inference())
    void forward_backward_encoded(cudaStream_t s, const __half* params, const float* coords, uint32_t n, const __half* dL_dout, __half* grads,
                                  cudaEvent_t after_mlp = nullptr, bool grid_grads_clean = false,
                                  const std::function<void(size_t, size_t)>* slice_ready = nullptr, float* dL_dcoords = nullptr);   // (offset, count) of `grads` complete on s   // recorded between the MLP and the hash-grid backward kernels
    const __half* last_inference_encoding() const { return enc_buf.as<__half>(); }
    __half* training_encoding(uint32_t n) { return enc_train_buf.get<__half>((size_t)n * dims.pos_width); }
};

struct EncodedMlp {
    GridLayout grid; uint32_t n_hidden = 2, n_output = 3;
    uint64_t off_mlp = 0, off_grid = 0, n_params = 0, n_matrix_params = 0;
    DeviceBuffer enc_buf, denc_buf, wgrad_buf;
    EncodedMlp(const Json& config, uint32_t n_pos_dims, uint32_t n_output_dims);
    void initialize_params(uint64_t seed, float* master_dev, cudaStream_t s);
    // x [n][x_stride] fp32; out [n][16] fp16
    void inference(cudaStream_t s, const __half* params, const float* x, uint32_t x_stride, uint32_t n, __half* out);
    void forward_backward(cudaStream_t s, const __half* params, const float* x, uint32_t x_stride, uint32_t n, const __half* dL_dout, __half* grads);
};

struct Optimizer {
    // Adam leaf
    float learning_rate = 1e-2f, beta1 = 0.9f, beta2 = 0.99f, epsilon = 1e-15f, l2_reg = 1e-6f;
    bool optimize_matrix = true, optimize_non_matrix = true;
    // ExponentialDecay
    uint32_t decay_start = 0, decay_interval = 0, decay_end = 0xFFFFFFFFu; float decay_base = 1.0f;
    // Ema
    float ema_decay = 0.0f;
    uint32_t step_count = 0; float current_lr = 1e-2f;
    explicit Optimizer(const Json& j);
    void update_hyperparams(const Json& j);   // testbed.cu:4339-4346
};

struct Trainer {
    uint64_t n_params = 0, n_matrix_params = 0, n_cap = 0;   // n_cap: allocated entries of params / grads / inference_params (multiple of 2048)
    float* master = nullptr; __half* params = nullptr; __half* inference_params = nullptr; __half* grads = nullptr;
    float* m = nullptr; float* v = nullptr; uint32_t* steps = nullptr; float* ema = nullptr;
    Optimizer opt;
    Trainer(uint64_t n_params, uint64_t n_matrix, const Json& optimizer_cfg);
    ~Trainer();
    Trainer(const Trainer&) = delete;
    void params_from_master(cudaStream_t s);             // fp16 copies after initialize_params
    void optimizer_step(cudaStream_t s, float loss_scale);
    void optimizer_step_shard(cudaStream_t s, float loss_scale, uint64_t start, uint64_t count);
    bool grid_grads_clean = false;   // the hash-table part of `grads` is all zero (left so by the last optimizer step)
    // north_star: "tables pinned in B200's L2 via access-policy windows when they fit".  params and grads are ONE allocation
    // [fp16 params | fp16 grads]; `s` gets a persisting access-policy window over it (hit ratio scaled to the carve-out when the
    // table is larger), so the encoder's gathers and scatters stay in L2 while Adam streams ~300 MB of fp32 state past them
    // (its loads / stores of that state carry evict-first hints).  Returns the bytes under the window (0: not applied).
    uint64_t apply_l2_policy(cudaStream_t s);
    uint64_t l2_window_bytes = 0; cudaStream_t l2_stream = nullptr; bool l2_tried = false;
};

}  // namespace ngp
