// Headless command-line driver over the C ABI — the role of the reference's src/main.cu (construct a Testbed, load a scene and a
// network config, train n steps, save a snapshot, render) with nothing but include/ngp_b200.h and the CUDA runtime: what a C++
// host application links against.  The scene is the synthetic NeRF-shape data set (no image decoder needed); a real scene goes
// through ngp_nerf_dataset_* instead.
//
//   ngp_train [--steps N] [--views N] [--res R] [--config network.json] [--snapshot out.ingp] [--load in.ingp] [--render out.ppm]
#include "../include/ngp_b200.h"
#include <cuda_runtime.h>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#define CHECK(call) do { if ((call) != 0) { std::fprintf(stderr, "%s failed: %s\n", #call, ngp_last_error()); return 1; } } while (0)
#define CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_)); return 1; } } while (0)

static const char* DEFAULT_CONFIG = R"({
  "loss": {"otype": "Huber"},
  "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
  "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819},
  "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
  "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]},
  "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}
})";

int main(int argc, char** argv) {
    int steps = 1000, n_views = 100, res = 400;
    std::string config = DEFAULT_CONFIG, snapshot_out, snapshot_in, render_out;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--steps") steps = std::atoi(next());
        else if (a == "--views") n_views = std::atoi(next());
        else if (a == "--res") res = std::atoi(next());
        else if (a == "--config") { std::ifstream f(next()); std::stringstream ss; ss << f.rdbuf(); config = ss.str(); }
        else if (a == "--snapshot") snapshot_out = next();
        else if (a == "--load") snapshot_in = next();
        else if (a == "--render") render_out = next();
        else { std::fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
    }
    cudaStream_t stream; CUDA(cudaStreamCreate(&stream));
    ngp_testbed_t tb; CHECK(ngp_testbed_create(/*Nerf*/ 0, &tb));

    // scene: n_views posed res x res RGBA8 views, rendered on the device and handed over as host buffers (like decoded image files)
    const float radius = 4.0f * 0.33f, focal = 0.5f * (float)res / std::tan(0.5f * 0.6911f);
    std::vector<std::vector<uint8_t>> pixels(n_views, std::vector<uint8_t>((size_t)res * res * 4));
    std::vector<ngp_training_image> views(n_views);
    void* dev = nullptr; CUDA(cudaMalloc(&dev, (size_t)res * res * 4));
    for (int v = 0; v < n_views; ++v) {
        CHECK(ngp_synth_render_view(stream, v, n_views, radius, focal, res, res, dev));
        CUDA(cudaMemcpyAsync(pixels[v].data(), dev, pixels[v].size(), cudaMemcpyDeviceToHost, stream));
        ngp_training_image& im = views[v];
        std::memset(&im, 0, sizeof(im));
        im.pixels = pixels[v].data(); im.data_type = 0; im.res[0] = im.res[1] = res; im.focal_length[0] = im.focal_length[1] = focal;
        im.principal_point[0] = im.principal_point[1] = 0.5f;
        CHECK(ngp_synth_camera(v, n_views, radius, im.xform));
    }
    CUDA(cudaStreamSynchronize(stream)); CUDA(cudaFree(dev));
    CHECK(ngp_testbed_set_nerf_dataset(tb, stream, n_views, views.data(), /*aabb_scale*/ 1));
    if (!snapshot_in.empty()) CHECK(ngp_testbed_load_snapshot(tb, stream, snapshot_in.c_str()));
    else CHECK(ngp_testbed_reload_network_from_json(tb, config.c_str()));
    std::printf("network: %llu parameters (%llu in the encoding)\n", (unsigned long long)ngp_testbed_n_params(tb), (unsigned long long)ngp_testbed_n_encoding_params(tb));

    ngp_train_stats st{};
    unsigned long long samples = 0;
    CUDA(cudaStreamSynchronize(stream));
    const auto t0 = std::chrono::steady_clock::now();
    for (int s = 0; s < steps; ++s) {
        CHECK(ngp_testbed_train(tb, stream, 1u << 18));
        CHECK(ngp_testbed_stats(tb, &st));
        samples += st.measured_batch_size;
        if ((s + 1) % 250 == 0 || s + 1 == steps) std::printf("step %u  loss %.6f  rays/batch %u  samples/step %u\n", st.training_step, st.loss, st.rays_per_batch, st.measured_batch_size);
    }
    CUDA(cudaStreamSynchronize(stream));
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (steps > 0) std::printf("%d steps in %.3f s: %.3f ms/step, %.1f M samples/s\n", steps, sec, sec / steps * 1e3, (double)samples / sec / 1e6);

    if (!snapshot_out.empty()) { CHECK(ngp_testbed_save_snapshot(tb, stream, snapshot_out.c_str(), 0, 1)); std::printf("saved %s\n", snapshot_out.c_str()); }
    if (!render_out.empty()) {   // view 0 again, linear premultiplied RGBA -> sRGB PPM (the headless stand-in for the window)
        std::vector<float> rgba((size_t)res * res * 4);
        const float fl[2] = {focal, focal}, sc[2] = {0.5f, 0.5f};
        CHECK(ngp_testbed_render(tb, stream, views[0].xform, fl, sc, res, res, /*spp*/ 4, 0, res, rgba.data(), /*out_is_device*/ 0, nullptr));
        std::ofstream f(render_out, std::ios::binary);
        f << "P6\n" << res << " " << res << "\n255\n";
        for (size_t p = 0; p < (size_t)res * res; ++p)
            for (int c = 0; c < 3; ++c) {
                const float l = rgba[p * 4 + c], s = l < 0.0031308f ? 12.92f * l : 1.055f * std::pow(l, 0.41666f) - 0.055f;
                f.put((char)(unsigned char)std::lround(std::fmin(std::fmax(s, 0.0f), 1.0f) * 255.0f));
            }
        std::printf("rendered %s\n", render_out.c_str());
    }
    ngp_testbed_destroy(tb);
    return 0;
}
