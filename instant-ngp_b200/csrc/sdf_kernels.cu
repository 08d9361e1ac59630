// SDF-primitive training kernels (see sdf_kernels.h).  `permute` and the element-wise losses live in tiny-cuda-nn, which is
// absent from the mount: they are restated from their published definitions (Kensler, "Correlated Multi-Jittered Sampling",
// Pixar TM 13-01, listing 4; tcnn losses: value = f(pred, target) / n_total, gradient = loss_scale * f' / n_total) — unpinned.
#include "sdf_kernels.h"

namespace ngp {

// Kensler's hash-based permutation of [0, l) keyed by p: a cycle-walking bijection on the next power of two
NGP_HD uint32_t kensler_permute(uint32_t i, uint32_t l, uint32_t p) {
    uint32_t w = l - 1;
    w |= w >> 1; w |= w >> 2; w |= w >> 4; w |= w >> 8; w |= w >> 16;
    do {
        i ^= p;             i *= 0xe170893du;
        i ^= p >> 16;
        i ^= (i & w) >> 4;
        i ^= p >> 8;        i *= 0x0929eb3fu;
        i ^= p >> 23;
        i ^= (i & w) >> 1;  i *= 1u | p >> 27;
                            i *= 0x6935fa69u;
        i ^= (i & w) >> 11; i *= 0x74dcb303u;
        i ^= (i & w) >> 2;  i *= 0x9e501cc3u;
        i ^= (i & w) >> 2;  i *= 0xc860a3dfu;
        i &= w;
        i ^= i >> 5;
    } while (i >= l);
    return (i + p) % l;
}

// one thread per output record: the 12-byte position and the distance of record permute(i) (testbed_sdf.cu:1298-1299 runs
// shuffle<vec3> and shuffle<float> with the same seed, i.e. the same permutation; one pass here)
__global__ void sdf_shuffle_kernel(uint32_t n_elements, uint32_t n_out, uint32_t seed, const float* __restrict__ pos, const float* __restrict__ dist,
                                   float* __restrict__ pos_out, float* __restrict__ dist_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    const uint32_t j = kensler_permute(i, n_elements, seed);
    const float x = __ldg(pos + (size_t)j * 3), y = __ldg(pos + (size_t)j * 3 + 1), z = __ldg(pos + (size_t)j * 3 + 2);
    pos_out[(size_t)i * 3] = x; pos_out[(size_t)i * 3 + 1] = y; pos_out[(size_t)i * 3 + 2] = z;
    dist_out[i] = __ldg(dist + j);
}

NGP_HD void tcnn_loss_element(int type, float prediction, float target, float n_total, float loss_scale, float& value, float& gradient) {
    const float difference = prediction - target;
    float g;
    switch (type) {
        case TLOSS_RELATIVE_L2: {
            const float denom = prediction * prediction + 0.01f;
            value = difference * difference / denom / n_total; g = 2.0f * difference / denom; break;
        }
        case TLOSS_L1: value = fabsf(difference) / n_total; g = copysignf(1.0f, difference); break;
        case TLOSS_MAPE: {
            const float scale = 1.0f / (fabsf(target) + 0.01f);
            value = fabsf(difference) * scale / n_total; g = scale * copysignf(1.0f, difference); break;
        }
        case TLOSS_SMAPE: {
            const float denom = 0.5f * (fabsf(prediction) + fabsf(target)) + 0.01f;
            value = fabsf(difference) / denom / n_total;
            g = copysignf(1.0f, difference) / denom; break;   // denominator held constant, like RelativeL2
        }
        default: value = difference * difference / n_total; g = 2.0f * difference; break;
    }
    gradient = loss_scale * g / n_total;
}

__global__ void tcnn_loss_kernel(int type, uint32_t n, uint32_t dims, const __half* __restrict__ pred, const float* __restrict__ target, float loss_scale,
                                 __half* __restrict__ dL, float* __restrict__ loss_values) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float n_total = (float)n * (float)dims;
    uint4 in[2];
    in[0] = reinterpret_cast<const uint4*>(pred + (size_t)i * 16)[0]; in[1] = reinterpret_cast<const uint4*>(pred + (size_t)i * 16)[1];
    const __half* p = reinterpret_cast<const __half*>(in);
    __align__(16) __half o[16];
    float l = 0.f;
#pragma unroll
    for (uint32_t c = 0; c < 16; ++c) {
        float v = 0.f, g = 0.f;
        if (c < dims) tcnn_loss_element(type, __half2float(p[c]), target[(size_t)i * dims + c], n_total, loss_scale, v, g);
        l += v; o[c] = __float2half_rn(g);
    }
    reinterpret_cast<uint4*>(dL + (size_t)i * 16)[0] = reinterpret_cast<uint4*>(o)[0];
    reinterpret_cast<uint4*>(dL + (size_t)i * 16)[1] = reinterpret_cast<uint4*>(o)[1];
    if (loss_values) loss_values[i] = l;
}

void sdf_shuffle(cudaStream_t s, uint32_t n_elements, uint32_t n_out, uint32_t seed, const float* positions, const float* distances,
                 float* positions_shuffled, float* distances_shuffled) {
    if (!n_out) return;
    NGP_PROFILE_SCOPE(s, "sdf_shuffle", n_out);
    sdf_shuffle_kernel<<<div_round_up(n_out, 256), 256, 0, s>>>(n_elements, n_out, seed, positions, distances, positions_shuffled, distances_shuffled);
    NGP_POST_LAUNCH();
}

void tcnn_loss(cudaStream_t s, int loss_type, uint32_t n, uint32_t dims, const __half* pred, const float* target, float loss_scale, __half* dL,
               float* loss_values) {
    NGP_CHECK(dims >= 1 && dims <= 16, "tcnn_loss: 1..16 output dims");
    NGP_PROFILE_SCOPE(s, "tcnn_loss", n);
    tcnn_loss_kernel<<<div_round_up(n, 256), 256, 0, s>>>(loss_type, n, dims, pred, target, loss_scale, dL, loss_values);
    NGP_POST_LAUNCH();
}

}  // namespace ngp
