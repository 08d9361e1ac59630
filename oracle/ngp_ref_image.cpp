// ORACLE — TEST INFRASTRUCTURE ONLY (see ngp_ref.h). Image primitive step restated from
// /root/reference/src/testbed_image.cu:62-76 (stratify2_kernel), :166-212 (eval_image_kernel_and_snap),
// :440-463 (image_coords_from_idx, image_mse_kernel).
#include "ngp_ref.h"
#include <cmath>
#include <algorithm>

static inline float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

extern "C" {
// positions [n][2] from pcg32(seed) advanced by skip: generate_random_uniform then stratify2_kernel
void ref_image_positions(uint64_t seed, uint64_t skip, uint32_t n, int stratify, float* pos) {
    ref_pcg32_fill_float(seed, skip, pos, (size_t)n * 2);
    uint32_t log2n = 0; while ((1u << log2n) < n) ++log2n;
    if (!stratify || (1u << log2n) != n || (log2n & 1)) return;
    uint32_t log2s = log2n / 2, size = 1u << log2s;
    for (uint32_t i = 0; i < n; ++i) {
        uint32_t ib = i & ((1u << log2n) - 1), x = ib & (size - 1), y = ib >> log2s;
        pos[i * 2] = pos[i * 2] / (float)size + ((float)x / (float)size);
        pos[i * 2 + 1] = pos[i * 2 + 1] / (float)size + ((float)y / (float)size);
    }
}
// low-discrepancy positions of image.random_mode Halton / Sobol: halton23_kernel (testbed_image.cu:41-46, halton<base> random_val.cuh:293-305) and
// sobol2_kernel (testbed_image.cu:48-53, ld_random_val_2d random_val.cuh:281-285); base_idx = batch_size * training_step
void ref_ld_random_val_2d(uint32_t index, uint32_t seed, float* out);   // ngp_ref_nerf.cpp
static float halton(size_t idx, unsigned base) {
    float f = 1, result = 0;
    while (idx > 0) { f /= (float)base; result += f * (float)(idx % base); idx /= base; }
    return result;
}
void ref_image_positions_ld(int mode, uint64_t base_idx, uint32_t seed, uint32_t n, float* pos) {
    for (uint32_t i = 0; i < n; ++i) {
        if (mode == 1) { pos[i * 2] = halton(base_idx + i, 2); pos[i * 2 + 1] = halton(base_idx + i, 3); }
        else ref_ld_random_val_2d((uint32_t)(base_idx + i), seed, pos + (size_t)i * 2);
    }
}
// image: RGBA float32 [H][W][4]; pos in/out (snapped to pixel centres when snap); target [n][3]
void ref_image_eval(const float* image, int W, int H, uint32_t n, float* pos, float* target, int snap, int linear_colors) {
    auto read_val = [&](int x, int y, float* v) {
        const float* p = image + ((size_t)y * W + x) * 4;
        for (int c = 0; c < 3; ++c) v[c] = linear_colors ? p[c] : linear_to_srgb(p[c]);
    };
    for (uint32_t i = 0; i < n; ++i) {
        float px = pos[i * 2], py = pos[i * 2 + 1], v[3];
        if (snap) {
            int ix = (int)floorf(px * (float)W), iy = (int)floorf(py * (float)H);
            pos[i * 2] = ((float)ix + 0.5f) / (float)W; pos[i * 2 + 1] = ((float)iy + 0.5f) / (float)H;
            ix = std::min(std::max(ix, 0), W - 1); iy = std::min(std::max(iy, 0), H - 1);
            read_val(ix, iy, v);
        } else {
            float fx = std::min(std::max(px * (float)W - 0.5f, 0.0f), (float)W - (1.0f + 1e-4f));
            float fy = std::min(std::max(py * (float)H - 0.5f, 0.0f), (float)H - (1.0f + 1e-4f));
            int ix = (int)fx, iy = (int)fy; float wx = fx - (float)ix, wy = fy - (float)iy;
            ix = std::min(std::max(ix, 0), W - 2); iy = std::min(std::max(iy, 0), H - 2);
            float a[3], b[3], c[3], d[3]; read_val(ix, iy, a); read_val(ix + 1, iy, b); read_val(ix, iy + 1, c); read_val(ix + 1, iy + 1, d);
            for (int k = 0; k < 3; ++k) v[k] = (1 - wx) * (1 - wy) * a[k] + wx * (1 - wy) * b[k] + (1 - wx) * wy * c[k] + wx * wy * d[k];
        }
        target[i * 3] = v[0]; target[i * 3 + 1] = v[1]; target[i * 3 + 2] = v[2];
    }
}
// tcnn L2 loss (create_loss "L2"; recalled, unpinned): per element diff^2/n_total ; grad = loss_scale*2*diff/n_total,
// n_total = batch*n_out(3). prediction fp16 [n][pred_stride]; dL fp16 [n][pred_stride] (padding cols zero)
double ref_l2_loss(const uint16_t* pred, uint32_t pred_stride, const float* target, uint32_t n, float loss_scale, uint16_t* dL) {
    double total = 0; const float n_total = (float)n * 3.0f;
    for (uint32_t i = 0; i < n; ++i) for (uint32_t c = 0; c < pred_stride; ++c) {
        if (c >= 3) { if (dL) dL[(size_t)i * pred_stride + c] = 0; continue; }
        _Float16 h; __builtin_memcpy(&h, &pred[(size_t)i * pred_stride + c], 2);
        float diff = (float)h - target[i * 3 + c];
        total += (double)(diff * diff / n_total);
        if (dL) { _Float16 g = (_Float16)(loss_scale * 2.0f * diff / n_total); __builtin_memcpy(&dL[(size_t)i * pred_stride + c], &g, 2); }
    }
    return total;
}
}
