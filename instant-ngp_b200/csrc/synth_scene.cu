// Synthetic-dataset renderer: ground-truth views of the analytic scene (synth_scene.h) written as RGBA8 into device
// memory.  Data synthesis for benchmarks / tests (BASELINE.json: NeRF-synthetic-shape scene), not part of the training path.
#include "synth_scene.h"
#include "../../include/ngp_b200.h"

namespace ngp {
__global__ void synth_view_kernel(uint32_t view, uint32_t n_views, float radius, float focal, int w, int h, float step, uint32_t* __restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    float xform[12]; synth_camera(view, n_views, radius, xform);
    out[(size_t)y * w + x] = synth_pixel(xform, focal, w, h, x, y, step);
}
}  // namespace ngp

extern "C" int ngp_synth_camera(uint32_t view, uint32_t n_views, float radius, float* xform_out) {
    if (!xform_out || n_views == 0) return 1;
    ngp::synth_camera(view, n_views, radius, xform_out);
    return 0;
}
extern "C" int ngp_synth_render_view(void* stream, uint32_t view, uint32_t n_views, float radius, float focal, int32_t w, int32_t h, void* dev_rgba8) {
    if (!dev_rgba8 || w <= 0 || h <= 0) return 1;
    const dim3 threads = {16, 8, 1}, blocks = {ngp::div_round_up((uint32_t)w, 16), ngp::div_round_up((uint32_t)h, 8), 1};
    ngp::synth_view_kernel<<<blocks, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(view, n_views, radius, focal, w, h, 1.73205080757f / 4096.f,
                                                                                          reinterpret_cast<uint32_t*>(dev_rgba8));
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
