// Image-primitive kernels (replaces src/testbed_image.cu:62-76 stratify2_kernel, :166-212 eval_image_kernel_and_snap,
// :440-463 image_coords_from_idx / image_mse_kernel) and the tcnn L2 loss used by Trainer::training_step
// (testbed_image.cu:276).  Sample generation, stratification and target lookup are fused into one kernel.
#include "image_kernels.h"
#include "nerf_device.cuh"

namespace ngp {

NGP_D float lin2srgb(float l) { return linear_to_srgb(l); }

NGP_D void image_lookup(const float* __restrict__ image, int W, int H, float& px, float& py, bool snap, bool linear_colors, float* v) {
    auto read_val = [&](int x, int y, float* o) {
        const float4 p = __ldg(reinterpret_cast<const float4*>(image) + (size_t)y * W + x);
        o[0] = linear_colors ? p.x : lin2srgb(p.x); o[1] = linear_colors ? p.y : lin2srgb(p.y); o[2] = linear_colors ? p.z : lin2srgb(p.z);
    };
    if (snap) {
        int ix = (int)floorf(px * (float)W), iy = (int)floorf(py * (float)H);
        px = ((float)ix + 0.5f) / (float)W; py = ((float)iy + 0.5f) / (float)H;
        ix = clampi(ix, 0, W - 1); iy = clampi(iy, 0, H - 1);
        read_val(ix, iy, v);
    } else {
        const float fx = fminf(fmaxf(px * (float)W - 0.5f, 0.0f), (float)W - (1.0f + 1e-4f));
        const float fy = fminf(fmaxf(py * (float)H - 0.5f, 0.0f), (float)H - (1.0f + 1e-4f));
        int ix = (int)fx, iy = (int)fy; const float wx = fx - (float)ix, wy = fy - (float)iy;
        ix = clampi(ix, 0, W - 2); iy = clampi(iy, 0, H - 2);
        float a[3], b[3], c[3], d[3]; read_val(ix, iy, a); read_val(ix + 1, iy, b); read_val(ix, iy + 1, c); read_val(ix + 1, iy + 1, d);
        for (int k = 0; k < 3; ++k) v[k] = (1 - wx) * (1 - wy) * a[k] + wx * (1 - wy) * b[k] + (1 - wx) * wy * c[k] + wx * wy * d[k];
    }
}

// halton<base>(idx) (random_val.cuh:293-305): radical inverse, accumulated in float exactly as the reference does
template <uint32_t BASE> NGP_D float halton(size_t idx) {
    float f = 1, result = 0;
    while (idx > 0) { f /= BASE; result += f * (float)(idx % BASE); idx /= BASE; }
    return result;
}

// positions by image.random_mode (testbed_image.cu:227-245): Halton -> halton23_kernel (:41-46), Sobol -> sobol2_kernel (:48-53, Owen-scrambled,
// seeded), otherwise element e of the [n][2] array is float #e of the pcg32 stream (generate_random_uniform), then stratify2_kernel (:62-76)
__global__ void image_batch_kernel(uint32_t n, Pcg32 rng, int random_mode, size_t base_idx, uint32_t seed, uint32_t log2_batch, bool stratify,
                                   const float* __restrict__ image, int W, int H, bool snap, bool linear_colors, float* __restrict__ pos, float* __restrict__ target) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float px, py;
    if (random_mode == IMG_RANDOM_HALTON) { px = halton<2>(base_idx + i); py = halton<3>(base_idx + i); }
    else if (random_mode == IMG_RANDOM_SOBOL) { float v[2]; ld_random_val_2d((uint32_t)(base_idx + i), seed, v); px = v[0]; py = v[1]; }
    else { rng.advance((int64_t)((uint64_t)i * 2)); px = rng.next_float(); py = rng.next_float(); }
    if (stratify) {
        const uint32_t log2s = log2_batch / 2, size = 1u << log2s;
        const uint32_t ib = i & ((1u << log2_batch) - 1), x = ib & (size - 1), y = ib >> log2s;
        px = px / (float)size + ((float)x / (float)size); py = py / (float)size + ((float)y / (float)size);
    }
    float v[3]; image_lookup(image, W, H, px, py, snap, linear_colors, v);
    pos[(size_t)i * 2] = px; pos[(size_t)i * 2 + 1] = py;
    target[(size_t)i * 3] = v[0]; target[(size_t)i * 3 + 1] = v[1]; target[(size_t)i * 3 + 2] = v[2];
}

__global__ void image_eval_coords_kernel(uint32_t n, uint32_t offset, const float* __restrict__ image, int W, int H, bool linear_colors,
                                         float* __restrict__ pos, float* __restrict__ target) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t idx = i + offset;
    const int x = clampi((int)(idx % (uint32_t)W), 0, W - 1), y = clampi((int)(idx / (uint32_t)W), 0, H - 1);
    float px = ((float)x + 0.5f) / (float)W, py = ((float)y + 0.5f) / (float)H;
    float v[3]; image_lookup(image, W, H, px, py, true, linear_colors, v);
    pos[(size_t)i * 2] = px; pos[(size_t)i * 2 + 1] = py;
    target[(size_t)i * 3] = v[0]; target[(size_t)i * 3 + 1] = v[1]; target[(size_t)i * 3 + 2] = v[2];
}

// image_mse_kernel (testbed_image.cu:452-463)
__global__ void image_se_kernel(uint32_t n, const float* __restrict__ target, const __half* __restrict__ pred, float* __restrict__ se, bool quantize) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float p = __half2float(pred[(size_t)i * 16 + c]);
        if (quantize) p = (float)clampi((int)(p * 255.0f + 0.5f), 0, 255) / 255.0f;
        const float d = target[(size_t)i * 3 + c] - p;
        acc += d * d;
    }
    se[i] = acc / 3.0f;
}

void image_generate_batch(cudaStream_t s, uint32_t n, Pcg32 rng, int random_mode, size_t base_idx, uint32_t seed, const float* image, int W, int H, bool snap,
                          bool linear_colors, float* pos, float* target) {
    uint32_t log2n = 0; while ((1u << log2n) < n) ++log2n;
    const bool can_stratify = random_mode == IMG_RANDOM_STRATIFIED && (1u << log2n) == n && (log2n % 2 == 0);
    image_batch_kernel<<<div_round_up(n, 256), 256, 0, s>>>(n, rng, random_mode, base_idx, seed, log2n, can_stratify, image, W, H, snap, linear_colors, pos, target);
    NGP_POST_LAUNCH();
}
void image_eval_coords(cudaStream_t s, uint32_t n, uint32_t offset, const float* image, int W, int H, bool linear_colors, float* pos, float* target) {
    image_eval_coords_kernel<<<div_round_up(n, 256), 256, 0, s>>>(n, offset, image, W, H, linear_colors, pos, target);
    NGP_POST_LAUNCH();
}
// from_rgba32<float> (common_device.cuh:771-798) as load_stbi_gpu calls it (common_device.cu:72: no white/black keying, no mask colour):
// sRGB bytes with straight alpha -> linear, alpha-premultiplied floats
__global__ void image_from_rgba8_kernel(uint64_t n, const uint32_t* __restrict__ px, float4* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t v = px[i];
    const float alpha = (float)(v >> 24) * (1.0f / 255.0f);
    out[i] = make_float4(srgb_to_linear((float)(v & 0xFF) * (1.0f / 255.0f)) * alpha, srgb_to_linear((float)((v >> 8) & 0xFF) * (1.0f / 255.0f)) * alpha,
                         srgb_to_linear((float)((v >> 16) & 0xFF) * (1.0f / 255.0f)) * alpha, alpha);
}
void image_from_rgba8(cudaStream_t s, uint64_t n, const uint8_t* rgba8, float* out) {
    image_from_rgba8_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, (const uint32_t*)rgba8, (float4*)out);
    NGP_POST_LAUNCH();
}
void image_squared_error(cudaStream_t s, uint32_t n, const float* target, const __half* pred, float* se, bool quantize) {
    image_se_kernel<<<div_round_up(n, 256), 256, 0, s>>>(n, target, pred, se, quantize);
    NGP_POST_LAUNCH();
}

}  // namespace ngp
