// ngp_nerf_network_inference_noalloc — NerfNetwork inference that ALLOCATES NOTHING (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).
//
// The regular entry point (ngp_nerf_network_inference, capi.cu) stages the hash-grid encoder's output — the [n x C] fp16 blk8
// tile array the fused MLP kernel pulls in with bulk-async copies — in a grow-on-demand cudaMalloc'ed buffer owned by the network
// object.  This variant runs the SAME two kernels (grid_forward -> nerf_mlp_forward) with that scratch in a file-scope
// `static __device__` array of fixed size: NGP_NOALLOC_MAX_SAMPLES (16384) samples x 32 columns x 2 B = 1 MiB, which every GPU
// that loads the library holds until the process exits.
//
// Rules the code below follows:
//   * in HOST code the array's name is a host-side stand-in, not a device address; the device address is obtained once per device,
//     with that device current, through cudaGetSymbolAddress (return code checked) and cached per device ordinal;
//   * sizes beyond the maximum are rejected with a clear error; no cudaMalloc / cudaFree / allocating library routine is called
//     (with per-kernel profiling, ngp_profile_enable, off — its event pool is the one thing that could create objects);
//   * every call shares the array and it keeps the previous call's contents.  Nothing has to start at zero: rows [0, n) of the
//     tile array are fully overwritten by grid_forward (padding columns included) before the MLP kernel reads exactly those rows,
//     and n is a whole number of 128-sample tiles.  Calls that overlap on the GPU (two streams, two host threads) RACE on the
//     array: serialise them.
#include "../../include/ngp_b200.h"
#include "network.h"
#include <mutex>

namespace ngp {
// defined in capi.cu: the network behind a handle, and the per-thread error string of the C ABI
NerfNetwork* capi_nerf_network(ngp_nerf_network_t h);
void capi_set_last_error(const char* msg);
}

namespace {

constexpr uint32_t MAX_SAMPLES = NGP_NOALLOC_MAX_SAMPLES;
constexpr uint32_t MAX_WIDTH = 32;                                   // padded encoder width of the NeRF path: 16 or 32
static __device__ __align__(256) __half g_enc_scratch[(size_t)MAX_SAMPLES * MAX_WIDTH];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
__half* g_scratch_ptr[MAX_DEVICES] = {};

// device address of the scratch array on the CURRENT device (looked up once per device)
__half* scratch_for_current_device() {
    int dev = -1;
    NGP_CUDA_CHECK(cudaGetDevice(&dev));
    NGP_CHECK(dev >= 0 && dev < MAX_DEVICES, "inference_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[dev]) {
        __half* p = nullptr;
        NGP_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_enc_scratch));
        NGP_CHECK(p != nullptr, "inference_noalloc: cudaGetSymbolAddress returned a null address");
        g_scratch_ptr[dev] = p;
    }
    return g_scratch_ptr[dev];
}

}  // namespace

extern "C" {

uint32_t ngp_noalloc_max_samples(void) { return MAX_SAMPLES; }

int ngp_nerf_network_inference_noalloc(ngp_nerf_network_t n, void* stream, const void* params, const void* coords, uint32_t count, void* out) {
    try {
        NGP_CHECK(n && params && coords && out, "null argument");
        int n_dev = 0;
        if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { cudaGetLastError(); throw ngp::Error("ngp_b200: no CUDA device available (this library has no CPU fallback)"); }
        NGP_CHECK(count <= MAX_SAMPLES, "inference_noalloc: count " + std::to_string(count) + " exceeds the static scratch capacity of " +
                                            std::to_string(MAX_SAMPLES) + " samples (NGP_NOALLOC_MAX_SAMPLES); split the batch or use ngp_nerf_network_inference");
        NGP_CHECK(count % ngp::TILE_ROWS == 0, "inference_noalloc: count must be a multiple of 128 (callers round to NGP_BATCH_GRANULARITY = 256)");
        if (count == 0) return 0;
        ngp::NerfNetwork* net = ngp::capi_nerf_network(n);
        NGP_CHECK(net->dims.pos_width <= MAX_WIDTH, "inference_noalloc: encoder width beyond the static scratch layout");
        __half* enc = scratch_for_current_device();
        cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
        const __half* p = reinterpret_cast<const __half*>(params);
        const float* c = reinterpret_cast<const float*>(coords);
        // the same two launches as NerfNetwork::inference (network.cu), scratch = the static array
        ngp::grid_forward(s, net->grid, p + net->off_grid, c, 7, count, enc, ngp::LAYOUT_BLK8);
        ngp::nerf_mlp_forward(s, net->dims, p + net->off_density, p + net->off_rgb, enc, c, count, reinterpret_cast<__half*>(out));
        return 0;
    } catch (const std::exception& e) { ngp::capi_set_last_error(e.what()); return 1; }
      catch (...) { ngp::capi_set_last_error("unknown error"); return 1; }
}

}  // extern "C"
