// ORACLE — TEST INFRASTRUCTURE ONLY (see ngp_ref.h).  SDF primitive training step with caller-supplied ground truth:
// /root/reference/src/python_api.cu:69-99 (override_sdf_training_data), src/testbed_sdf.cu:1289-1311 (train_sdf).
// shuffle<T>/permute and the element-wise losses are tiny-cuda-nn code (absent from the mount): restated from the published
// definitions (Kensler, "Correlated Multi-Jittered Sampling", Pixar TM 13-01; tcnn loss = f / n_total) — PARITY UNPINNED.
#include "ngp_ref.h"
#include <cmath>
#include <cstring>

static uint32_t permute(uint32_t i, uint32_t l, uint32_t p) {
    uint32_t w = l - 1;
    w |= w >> 1; w |= w >> 2; w |= w >> 4; w |= w >> 8; w |= w >> 16;
    do {
        i ^= p; i *= 0xe170893du;
        i ^= p >> 16;
        i ^= (i & w) >> 4;
        i ^= p >> 8; i *= 0x0929eb3fu;
        i ^= p >> 23;
        i ^= (i & w) >> 1; i *= 1u | p >> 27;
        i *= 0x6935fa69u;
        i ^= (i & w) >> 11; i *= 0x74dcb303u;
        i ^= (i & w) >> 2; i *= 0x9e501cc3u;
        i ^= (i & w) >> 2; i *= 0xc860a3dfu;
        i &= w;
        i ^= i >> 5;
    } while (i >= l);
    return (i + p) % l;
}

extern "C" {
uint32_t ref_permute(uint32_t i, uint32_t l, uint32_t p) { return permute(i, l, p); }

// python_api.cu:81-90: mesh space -> unit cube
void ref_sdf_normalize(uint32_t n, const float* points, const float* distances, const float* raw_min, const float* raw_max, float mesh_scale,
                       float* points_out, float* distances_out) {
    for (uint32_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k)
            points_out[(size_t)i * 3 + k] = (points[(size_t)i * 3 + k] - raw_min[k]) / mesh_scale + 0.5f * (1.0f - (raw_max[k] - raw_min[k]) / mesh_scale);
        distances_out[i] = distances[i] / mesh_scale;
    }
}

// testbed_sdf.cu:1298-1299: shuffle<vec3> and shuffle<float> (stride 1) with seed = training step; first n_out records
void ref_sdf_shuffle(uint32_t n_elements, uint32_t n_out, uint32_t seed, const float* pos, const float* dist, float* pos_out, float* dist_out) {
    for (uint32_t i = 0; i < n_out; ++i) {
        const uint32_t j = permute(i, n_elements, seed);
        memcpy(pos_out + (size_t)i * 3, pos + (size_t)j * 3, 12);
        dist_out[i] = dist[j];
    }
}

// element-wise tcnn loss: type 0 L2, 1 RelativeL2, 2 L1, 3 MAPE, 4 SMAPE.  pred fp16 [n][pred_stride], target [n][dims];
// dL fp16 [n][pred_stride] (padding zero).  The relative losses (RelativeL2, SMAPE) treat their normaliser as a constant in the
// gradient.  Returns the summed loss (Trainer::loss = reduce_sum of the values)
double ref_tcnn_loss(int type, const uint16_t* pred, uint32_t pred_stride, const float* target, uint32_t n, uint32_t dims, float loss_scale, uint16_t* dL) {
    double total = 0; const float n_total = (float)n * (float)dims;
    for (uint32_t i = 0; i < n; ++i) for (uint32_t c = 0; c < pred_stride; ++c) {
        if (c >= dims) { if (dL) dL[(size_t)i * pred_stride + c] = 0; continue; }
        _Float16 h; memcpy(&h, &pred[(size_t)i * pred_stride + c], 2);
        const float prediction = (float)h, t = target[(size_t)i * dims + c], difference = prediction - t;
        float value, g;
        switch (type) {
            case 1: { const float d = prediction * prediction + 0.01f; value = difference * difference / d / n_total; g = 2.0f * difference / d; break; }
            case 2: value = fabsf(difference) / n_total; g = copysignf(1.0f, difference); break;
            case 3: { const float s = 1.0f / (fabsf(t) + 0.01f); value = fabsf(difference) * s / n_total; g = s * copysignf(1.0f, difference); break; }
            case 4: {
                const float d = 0.5f * (fabsf(prediction) + fabsf(t)) + 0.01f;
                value = fabsf(difference) / d / n_total;
                g = copysignf(1.0f, difference) / d; break;
            }
            default: value = difference * difference / n_total; g = 2.0f * difference; break;
        }
        total += (double)value;
        if (dL) { _Float16 gh = (_Float16)(loss_scale * g / n_total); memcpy(&dL[(size_t)i * pred_stride + c], &gh, 2); }
    }
    return total;
}
}
