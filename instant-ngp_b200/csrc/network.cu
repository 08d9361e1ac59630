// Host-side model objects (see network.h).  Parameter layout follows NerfNetwork::set_params_impl
// (nerf_network.h:430-457): [density MLP | rgb MLP | position encoding | direction encoding (0 params)].
#include "network.h"
#include <cmath>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <algorithm>

namespace ngp {

MlpConfig parse_mlp_config(const Json& j, const char* what) {
    MlpConfig c;
    const std::string otype = to_lower(j.value("otype", "FullyFusedMLP"));
    NGP_CHECK(otype == "fullyfusedmlp" || otype == "megakernelmlp", std::string(what) + ": only otype FullyFusedMLP is implemented on this path (got '" + otype + "')");
    NGP_CHECK(to_lower(j.value("activation", "ReLU")) == "relu", std::string(what) + ": only ReLU hidden activation is implemented");
    NGP_CHECK(to_lower(j.value("output_activation", "None")) == "none", std::string(what) + ": only output_activation None is implemented");
    c.n_neurons = (uint32_t)j.value("n_neurons", 64.0);
    c.n_hidden_layers = (uint32_t)j.value("n_hidden_layers", 2.0);
    c.n_output_dims = (uint32_t)j.value("n_output_dims", 16.0);
    NGP_CHECK(c.n_neurons == 64, std::string(what) + ": n_neurons must be 64 on the tcgen05 path");
    NGP_CHECK(c.n_hidden_layers >= 1 && c.n_hidden_layers <= 3, std::string(what) + ": n_hidden_layers must be 1..3");
    return c;
}

// encoding config derivation: testbed.cu:3961-4005 (fork: per_level_scale is never injected -> tcnn default 2.0)
void grid_from_json(GridLayout& g, const Json& enc, uint32_t n_pos_dims) {
    const std::string otype = to_lower(enc.value("otype", "HashGrid"));
    NGP_CHECK(otype.find("grid") != std::string::npos && otype != "densegrid" && otype != "tiledgrid",
              "encoding: only otype HashGrid is implemented on this path (got '" + otype + "')");
    const uint32_t F = (uint32_t)enc.value("n_features_per_level", 2.0);
    uint32_t L = (uint32_t)enc.value("n_levels", 16.0);
    if (enc.contains("n_features") && enc.value("n_features", 0.0) > 0) L = (uint32_t)enc.value("n_features", 0.0) / F;
    const uint32_t log2T = (uint32_t)enc.value("log2_hashmap_size", 19.0);
    uint32_t base = (uint32_t)enc.value("base_resolution", 0.0);
    if (!base) base = 1u << ((uint32_t)enc.value("log2_hashmap_size", 15.0) / n_pos_dims);
    const float b = (float)enc.value("per_level_scale", 2.0);
    grid_make_layout(g, n_pos_dims, L, F, log2T, base, b);
}

void mlp_init_host(uint32_t in_w, uint32_t n_hidden, Pcg32& rng, float* p) {
    uint32_t cw = in_w;
    for (uint32_t l = 0; l <= n_hidden; ++l) {
        const uint32_t rows = (l == n_hidden) ? MLP_OUT_PAD : MLP_WIDTH;
        const float s = sqrtf(6.0f / (float)(rows + cw));
        for (uint32_t i = 0; i < rows * cw; ++i) *p++ = (rng.next_float() * 2.0f - 1.0f) * s;
        cw = MLP_WIDTH;
    }
}

static void check_dir_encoding(const Json& cfg) {
    if (!cfg.contains("dir_encoding")) return;   // reference default is the SH composite as well
    const Json& d = cfg.at("dir_encoding");
    const Json* sh = &d;
    if (to_lower(d.value("otype", "")) == "composite") {
        NGP_CHECK(d.contains("nested") && !d.at("nested").arr.empty(), "dir_encoding: Composite needs nested encodings");
        sh = &d.at("nested").arr[0];
        for (size_t i = 1; i < d.at("nested").arr.size(); ++i)
            NGP_CHECK(to_lower(d.at("nested").arr[i].value("otype", "")) == "identity", "dir_encoding: nested[1..] must be Identity");
    }
    NGP_CHECK(to_lower(sh->value("otype", "")) == "sphericalharmonics" && (int)sh->value("degree", 4.0) == 4,
              "dir_encoding: only SphericalHarmonics degree 4 (+Identity) is implemented on this path");
}

// ---------------------------------------------------------------------------------------------------------------
NerfNetwork::NerfNetwork(const Json& cfg) {
    NGP_CHECK(cfg.contains("encoding") && cfg.contains("network"), "network config needs 'encoding' and 'network'");
    grid_from_json(grid, cfg.at("encoding"), 3);
    check_dir_encoding(cfg);
    MlpConfig dn = parse_mlp_config(cfg.at("network"), "network");
    MlpConfig rn = parse_mlp_config(cfg.contains("rgb_network") ? cfg.at("rgb_network") : cfg.at("network"), "rgb_network");
    NGP_CHECK(!cfg.at("network").contains("n_output_dims") || dn.n_output_dims == 16, "network.n_output_dims must be 16");
    NGP_CHECK(dn.n_hidden_layers <= 2 && rn.n_hidden_layers <= 2, "NeRF path: density/rgb networks support 1..2 hidden layers");
    dims.pos_width = grid.padded_width;
    NGP_CHECK(dims.pos_width == 16 || dims.pos_width == 32, "NeRF path: n_levels*n_features_per_level must pad to 16 or 32");
    dims.density_hidden = dn.n_hidden_layers; dims.rgb_hidden = rn.n_hidden_layers;
    const uint64_t nd = mlp_n_params(dims.pos_width, dims.density_hidden), nr = mlp_n_params(32, dims.rgb_hidden);
    off_density = 0; off_rgb = nd; off_grid = nd + nr;
    n_matrix_params = nd + nr;
    n_params = n_matrix_params + grid.n_params;
}

void NerfNetwork::initialize_params(uint64_t seed, float* master_dev, cudaStream_t s) {
    // initialize_params order (nerf_network.h:445-457): density MLP, rgb MLP, pos encoding; seed testbed.h:609
    Pcg32 rng(seed);
    std::vector<float> host(n_matrix_params);
    mlp_init_host(dims.pos_width, dims.density_hidden, rng, host.data() + off_density);
    mlp_init_host(32, dims.rgb_hidden, rng, host.data() + off_rgb);
    NGP_CUDA_CHECK(cudaMemcpyAsync(master_dev, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    grid_init(s, grid, rng, master_dev + off_grid);
}

void NerfNetwork::inference(cudaStream_t s, const __half* params, const float* coords, uint32_t n, __half* out, const uint32_t* n_dyn) {
    __half* enc = enc_buf.get<__half>((size_t)n * dims.pos_width);
    grid_forward(s, grid, params + off_grid, coords, 7, n, enc, LAYOUT_BLK8, n_dyn, max_level_gpu);
    nerf_mlp_forward(s, dims, params + off_density, params + off_rgb, enc, coords, n, out, n_dyn);
}

void NerfNetwork::density(cudaStream_t s, const __half* params, const float* pos, uint32_t pos_stride, uint32_t n, __half* out) {
    __half* enc = enc_buf.get<__half>((size_t)n * dims.pos_width);
    grid_forward(s, grid, params + off_grid, pos, pos_stride, n, enc, LAYOUT_BLK8);
    nerf_mlp_density(s, dims, params + off_density, enc, n, out);
}

void NerfNetwork::forward_backward(cudaStream_t s, const __half* params, const float* coords, uint32_t n, const __half* dL_dout, __half* grads, float* dL_dcoords) {
    __half* enc = enc_buf.get<__half>((size_t)n * dims.pos_width);
    __half* denc = denc_buf.get<__half>((size_t)n * dims.pos_width);
    float* wg = wgrad_buf.get<float>(n_matrix_params);
    NGP_CUDA_CHECK(cudaMemsetAsync(wg, 0, n_matrix_params * sizeof(float), s));
    NGP_CUDA_CHECK(cudaMemsetAsync(grads + off_grid, 0, grid.n_params * sizeof(__half), s));
    grid_forward(s, grid, params + off_grid, coords, 7, n, enc, LAYOUT_BLK8, nullptr, max_level_gpu);
    nerf_mlp_backward(s, dims, params + off_density, params + off_rgb, enc, coords, n, dL_dout, denc, wg, dL_dcoords);
    cast_f32_to_half(s, wg, grads, n_matrix_params);
    grid_backward(s, grid, coords, 7, n, denc, LAYOUT_BLK8, grads + off_grid, max_level_gpu);
    if (dL_dcoords) grid_backward_input(s, grid, params + off_grid, coords, 7, n, denc, LAYOUT_BLK8, dL_dcoords, 7, max_level_gpu);
}

void NerfNetwork::forward_backward_encoded(cudaStream_t s, const __half* params, const float* coords, uint32_t n, const __half* dL_dout, __half* grads,
                                           cudaEvent_t after_mlp, bool grid_grads_clean, const std::function<void(size_t, size_t)>* slice_ready, float* dL_dcoords) {
    const __half* enc = enc_train_buf.as<__half>();
    __half* denc = denc_buf.get<__half>((size_t)n * dims.pos_width);
    float* wg = wgrad_buf.get<float>(n_matrix_params);
    NGP_CUDA_CHECK(cudaMemsetAsync(wg, 0, n_matrix_params * sizeof(float), s));
    if (!grid_grads_clean) NGP_CUDA_CHECK(cudaMemsetAsync(grads + off_grid, 0, grid.n_params * sizeof(__half), s));
    nerf_mlp_backward(s, dims, params + off_density, params + off_rgb, enc, coords, n, dL_dout, denc, wg, dL_dcoords);
    cast_f32_to_half(s, wg, grads, n_matrix_params);
    if (after_mlp) NGP_CUDA_CHECK(cudaEventRecord(after_mlp, s));
    if (dL_dcoords) grid_backward_input(s, grid, params + off_grid, coords, 7, n, denc, LAYOUT_BLK8, dL_dcoords, 7, max_level_gpu);
    if (!slice_ready) { grid_backward(s, grid, coords, 7, n, denc, LAYOUT_BLK8, grads + off_grid, max_level_gpu); return; }
    // data-parallel step: hand every finished slice of the gradient vector to the caller (who all-reduces it on another stream).
    // The hash-grid pass can be cut into `groups` launches of consecutive level chunks (NGP_DP_SLICES) so that the collective
    // starts under the later chunks.  Measured on 8 B200 (24.4 MB table, all-reduce 0.125 ms): 1 group 0.616 ms/step, 2 groups
    // 0.653, 4 groups 0.728 — each extra collective costs more in launch latency and SMs taken from the scatter kernels than
    // the overlap returns, so the default is ONE all-reduce of the whole vector behind the backward pass.
    const uint32_t levels_per_chunk = 8 / grid.n_features, n_chunks = div_round_up(grid.n_levels, levels_per_chunk);
    uint32_t groups = 1;
    if (const char* e = getenv("NGP_DP_SLICES")) groups = (uint32_t)std::max(1, atoi(e));
    groups = std::min(groups, n_chunks);
    bool mlp_sent = false;
    for (uint32_t gi = 0; gi < groups; ++gi) {
        const uint32_t c0 = gi * n_chunks / groups, c1 = (gi + 1) * n_chunks / groups;
        const uint32_t l0 = c0 * levels_per_chunk, l1 = std::min(c1 * levels_per_chunk, grid.n_levels);
        grid_backward_chunks(s, grid, coords, 7, n, denc, LAYOUT_BLK8, grads + off_grid, c0, c1 - c0, max_level_gpu);
        // the MLP weights (first in the vector) travel with the first slice when it starts at level 0: one collective fewer
        const size_t off = off_grid + (size_t)grid.offsets[l0] * grid.n_features, cnt = (size_t)(grid.offsets[l1] - grid.offsets[l0]) * grid.n_features;
        if (gi == 0 && off_grid == n_matrix_params) { (*slice_ready)(0, n_matrix_params + cnt); mlp_sent = true; }
        else (*slice_ready)(off, cnt);
    }
    if (!mlp_sent) (*slice_ready)(0, n_matrix_params);
}

// ---------------------------------------------------------------------------------------------------------------
EncodedMlp::EncodedMlp(const Json& cfg, uint32_t n_pos_dims, uint32_t n_output_dims) {
    NGP_CHECK(cfg.contains("encoding") && cfg.contains("network"), "network config needs 'encoding' and 'network'");
    grid_from_json(grid, cfg.at("encoding"), n_pos_dims);
    MlpConfig mc = parse_mlp_config(cfg.at("network"), "network");
    NGP_CHECK(grid.padded_width == 16 || grid.padded_width == 32, "n_levels*n_features_per_level must pad to 16 or 32");
    NGP_CHECK(n_output_dims <= 16, "n_output_dims must be <= 16");
    n_hidden = mc.n_hidden_layers; n_output = n_output_dims;
    off_mlp = 0; n_matrix_params = mlp_n_params(grid.padded_width, n_hidden); off_grid = n_matrix_params;
    n_params = n_matrix_params + grid.n_params;
}

void EncodedMlp::initialize_params(uint64_t seed, float* master_dev, cudaStream_t s) {
    Pcg32 rng(seed);
    std::vector<float> host(n_matrix_params);
    mlp_init_host(grid.padded_width, n_hidden, rng, host.data());
    NGP_CUDA_CHECK(cudaMemcpyAsync(master_dev, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    grid_init(s, grid, rng, master_dev + off_grid);
}

void EncodedMlp::inference(cudaStream_t s, const __half* params, const float* x, uint32_t x_stride, uint32_t n, __half* out) {
    __half* enc = enc_buf.get<__half>((size_t)n * grid.padded_width);
    grid_forward(s, grid, params + off_grid, x, x_stride, n, enc, LAYOUT_BLK8);
    mlp_forward(s, grid.padded_width, n_hidden, params + off_mlp, enc, n, out, 16);
}

void EncodedMlp::forward_backward(cudaStream_t s, const __half* params, const float* x, uint32_t x_stride, uint32_t n, const __half* dL_dout, __half* grads) {
    __half* enc = enc_buf.get<__half>((size_t)n * grid.padded_width);
    __half* denc = denc_buf.get<__half>((size_t)n * grid.padded_width);
    float* wg = wgrad_buf.get<float>(n_matrix_params);
    NGP_CUDA_CHECK(cudaMemsetAsync(wg, 0, n_matrix_params * sizeof(float), s));
    NGP_CUDA_CHECK(cudaMemsetAsync(grads + off_grid, 0, grid.n_params * sizeof(__half), s));
    grid_forward(s, grid, params + off_grid, x, x_stride, n, enc, LAYOUT_BLK8);
    mlp_backward(s, grid.padded_width, n_hidden, params + off_mlp, enc, n, dL_dout, 16, denc, wg);
    cast_f32_to_half(s, wg, grads, n_matrix_params);
    grid_backward(s, grid, x, x_stride, n, denc, LAYOUT_BLK8, grads + off_grid);
}

// ---------------------------------------------------------------------------------------------------------------
Optimizer::Optimizer(const Json& j) { update_hyperparams(j); current_lr = learning_rate; }

void Optimizer::update_hyperparams(const Json& j) {
    const Json* cur = &j;
    for (int depth = 0; depth < 8; ++depth) {
        const std::string otype = to_lower(cur->value("otype", "Adam"));
        if (otype == "ema") ema_decay = (float)cur->value("decay", 0.99);
        else if (otype == "exponentialdecay") {
            decay_start = (uint32_t)cur->value("decay_start", 0.0); decay_interval = (uint32_t)cur->value("decay_interval", 0.0);
            decay_base = (float)cur->value("decay_base", 1.0);
            decay_end = cur->contains("decay_end") ? (uint32_t)cur->value("decay_end", 0.0) : 0xFFFFFFFFu;
        } else if (otype == "adam") {
            learning_rate = (float)cur->value("learning_rate", 1e-3); beta1 = (float)cur->value("beta1", 0.9); beta2 = (float)cur->value("beta2", 0.999);
            epsilon = (float)cur->value("epsilon", 1e-8); l2_reg = (float)cur->value("l2_reg", 1e-8);
            optimize_matrix = cur->value("optimize_matrix_params", true); optimize_non_matrix = cur->value("optimize_non_matrix_params", true);
        } else throw Error("optimizer: unsupported otype '" + otype + "' (Ema / ExponentialDecay / Adam are implemented)");
        if (!cur->contains("nested")) break;
        cur = &cur->at("nested");
    }
}

Trainer::Trainer(uint64_t n, uint64_t n_matrix, const Json& optimizer_cfg) : n_params(n), n_matrix_params(n_matrix), opt(optimizer_cfg) {
    // capacity: a multiple of 2048 entries, so that the vector splits into 1 / 2 / 4 / 8 / 16 equal, 128-entry-aligned shards for the
    // sharded optimizer (reduce-scatter / all-gather work in place over [0, n_cap)); grads follow params in the same allocation
    n_cap = next_multiple(n, 2048);
    NGP_CUDA_CHECK(cudaMalloc(&master, n * 4)); NGP_CUDA_CHECK(cudaMalloc(&params, n_cap * 2 * 2)); grads = params + n_cap;
    NGP_CUDA_CHECK(cudaMemset(params, 0, n_cap * 2 * 2));
    NGP_CUDA_CHECK(cudaMalloc(&m, n * 4)); NGP_CUDA_CHECK(cudaMalloc(&v, n * 4)); NGP_CUDA_CHECK(cudaMalloc(&steps, n * 4));
    NGP_CUDA_CHECK(cudaMemset(m, 0, n * 4)); NGP_CUDA_CHECK(cudaMemset(v, 0, n * 4)); NGP_CUDA_CHECK(cudaMemset(steps, 0, n * 4));
    NGP_CUDA_CHECK(cudaMemset(grads, 0, n * 2));
    if (opt.ema_decay > 0.0f) {
        NGP_CUDA_CHECK(cudaMalloc(&ema, n * 4)); NGP_CUDA_CHECK(cudaMemset(ema, 0, n * 4));
        NGP_CUDA_CHECK(cudaMalloc(&inference_params, n_cap * 2)); NGP_CUDA_CHECK(cudaMemset(inference_params, 0, n_cap * 2));
    } else inference_params = params;
}

Trainer::~Trainer() {
    if (inference_params && inference_params != params) cudaFree(inference_params);
    cudaFree(master); cudaFree(params); cudaFree(m); cudaFree(v); cudaFree(steps); if (ema) cudaFree(ema);
}

uint64_t Trainer::apply_l2_policy(cudaStream_t s) {
    if (l2_tried && l2_stream == s) return l2_window_bytes;
    l2_tried = true; l2_stream = s; l2_window_bytes = 0;
    // OFF unless NGP_L2_PERSIST=1.  Measured on B200 at the headline config (profiles/r02_ab_l2_red.json): with the 48.8 MB window
    // grid_forward is unchanged (0.115 ms: its gathers already hit — the bound is the L2 sector rate, not DRAM misses), grid_backward
    // gains 1 % and Adam LOSES 24 % (0.074 -> 0.091 ms: its fp16 reads/writes of the table become persisting accesses and the
    // carve-out takes 49 of 126 MB from the fp32 state streaming through) — 0.552 -> 0.562 ms per step overall.
    { const char* e = getenv("NGP_L2_PERSIST"); if (!e || atoi(e) == 0) return 0; }
    int dev = 0, max_persist = 0, max_window = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
    if (max_persist <= 0 || max_window <= 0) return 0;
    const uint64_t want = (uint64_t)((const char*)(grads + n_params) - (const char*)params);
    const uint64_t window = std::min<uint64_t>(want, (uint64_t)max_window), carve = std::min<uint64_t>(window, (uint64_t)max_persist);
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) != cudaSuccess) { cudaGetLastError(); return 0; }
    cudaStreamAttrValue attr{};
    attr.accessPolicyWindow.base_ptr = params; attr.accessPolicyWindow.num_bytes = window;
    attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)carve / (double)window);
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting; attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    if (cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr) != cudaSuccess) { cudaGetLastError(); return 0; }
    l2_window_bytes = window;
    if (getenv("NGP_DEBUG")) fprintf(stderr, "[ngp] L2 policy: window %llu B, carve-out %llu B (max persisting %d, max window %d), hit ratio %.2f\n",
                                     (unsigned long long)window, (unsigned long long)carve, max_persist, max_window, attr.accessPolicyWindow.hitRatio);
    return window;
}

void Trainer::params_from_master(cudaStream_t s) {
    cast_f32_to_half(s, master, params, n_params);
    if (inference_params != params) NGP_CUDA_CHECK(cudaMemcpyAsync(inference_params, params, n_params * 2, cudaMemcpyDeviceToDevice, s));
}

// Sharded step (data parallel): the caller has reduce-scattered the gradients, so entries [start, start + count) of
// `grads` hold the global sums; Adam (+EMA) runs on that slice only — 1/world of the HBM traffic — and the caller all-gathers the
// fp16 parameters afterwards.  Optimizer state outside the slice is never touched on this rank (its owner keeps it).
void Trainer::optimizer_step_shard(cudaStream_t s, float loss_scale, uint64_t start, uint64_t count) {
    NGP_CHECK(start % 8 == 0, "optimizer_step_shard: shard start must be a multiple of 8");
    const uint64_t end = std::min(start + count, n_params);
    AdamHyper h{};
    h.learning_rate = opt.current_lr; h.beta1 = opt.beta1; h.beta2 = opt.beta2; h.epsilon = opt.epsilon; h.l2_reg = opt.l2_reg;
    h.loss_scale = loss_scale;
    h.n_matrix_params = n_matrix_params > start ? std::min(n_matrix_params - start, end > start ? end - start : 0) : 0;
    h.optimize_matrix = opt.optimize_matrix; h.optimize_non_matrix = opt.optimize_non_matrix;
    h.ema_decay = opt.ema_decay; h.ema_step = opt.step_count + 1;
    h.clear_consumed_grads = 0; grid_grads_clean = false;      // the other ranks' slices hold partial sums: the whole table is cleared before the next pass
    if (end > start)
        adam_step(s, h, end - start, master + start, params + start, grads + start, m + start, v + start, steps + start, ema ? ema + start : nullptr,
                  ema ? inference_params + start : nullptr);
    ++opt.step_count;
    const uint32_t sc = opt.step_count;
    if (opt.decay_interval > 0 && sc >= opt.decay_start && sc < opt.decay_end && (sc - opt.decay_start) % opt.decay_interval == 0) opt.current_lr *= opt.decay_base;
}

void Trainer::optimizer_step(cudaStream_t s, float loss_scale) {
    AdamHyper h{};
    h.learning_rate = opt.current_lr; h.beta1 = opt.beta1; h.beta2 = opt.beta2; h.epsilon = opt.epsilon; h.l2_reg = opt.l2_reg;
    h.loss_scale = loss_scale; h.n_matrix_params = n_matrix_params;
    h.optimize_matrix = opt.optimize_matrix; h.optimize_non_matrix = opt.optimize_non_matrix;
    h.ema_decay = opt.ema_decay; h.ema_step = opt.step_count + 1;
    h.clear_consumed_grads = 1; grid_grads_clean = true;
    apply_l2_policy(s);
    adam_step(s, h, n_params, master, params, grads, m, v, steps, ema, ema ? inference_params : nullptr);
    ++opt.step_count;
    const uint32_t sc = opt.step_count;
    if (opt.decay_interval > 0 && sc >= opt.decay_start && sc < opt.decay_end && (sc - opt.decay_start) % opt.decay_interval == 0) opt.current_lr *= opt.decay_base;
}

}  // namespace ngp
