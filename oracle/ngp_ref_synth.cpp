// ORACLE — TEST INFRASTRUCTURE ONLY.  CPU copy of the procedural analytic scene used for the synthetic NeRF dataset
// (BASELINE.json; generator spec SURVEY.md §8d config 4).  No reference counterpart (the reference loads images).
#include <cmath>
#include <cstdint>
#include <algorithm>
extern "C" {
static inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
static inline float smooth(float x) { x = clampf(x, 0.f, 1.f); return x * x * (3.f - 2.f * x); }
struct Prim { float c[3]; float r[3]; float rgb[3]; int box; };
static const Prim PRIMS[4] = {
    {{0.50f, 0.47f, 0.50f}, {0.17f, 0.17f, 0.17f}, {0.90f, 0.30f, 0.20f}, 0},
    {{0.68f, 0.62f, 0.42f}, {0.10f, 0.10f, 0.10f}, {0.20f, 0.80f, 0.30f}, 0},
    {{0.34f, 0.64f, 0.60f}, {0.08f, 0.08f, 0.08f}, {0.20f, 0.40f, 0.90f}, 0},
    {{0.50f, 0.26f, 0.50f}, {0.27f, 0.03f, 0.27f}, {0.85f, 0.80f, 0.65f}, 1},
};
static void field(const float* p, const float* dir, float* sigma, float* rgb) {
    *sigma = 0.f; rgb[0] = rgb[1] = rgb[2] = 0.f;
    for (int k = 0; k < 4; ++k) {
        const Prim& q = PRIMS[k]; float sd;
        if (q.box) sd = fmaxf(fmaxf(fabsf(p[0] - q.c[0]) - q.r[0], fabsf(p[1] - q.c[1]) - q.r[1]), fabsf(p[2] - q.c[2]) - q.r[2]);
        else { float dx = p[0] - q.c[0], dy = p[1] - q.c[1], dz = p[2] - q.c[2]; sd = sqrtf(dx * dx + dy * dy + dz * dz) - q.r[0]; }
        float s = 400.f * smooth(0.5f - sd / 0.02f);
        if (s > 0.f) {
            float l = dir[0] * 0.3f + dir[1] * 0.8f + dir[2] * 0.5f, vd = 0.8f + 0.2f * l * l;
            for (int c = 0; c < 3; ++c) rgb[c] += s * vd * q.rgb[c];
            *sigma += s;
        }
    }
    if (*sigma > 0.f) for (int c = 0; c < 3; ++c) rgb[c] /= *sigma;
}
void ref_synth_camera(uint32_t i, uint32_t n, float radius, float* x) {
    const float golden = 2.39996322972865332f;
    float h = 0.08f + 0.87f * ((float)i + 0.5f) / (float)n, r = sqrtf(fmaxf(1.f - h * h, 0.f)), phi = golden * (float)i;
    float pos[3] = {0.5f + radius * r * cosf(phi), 0.5f + radius * h, 0.5f + radius * r * sinf(phi)};
    float f[3] = {0.5f - pos[0], 0.5f - pos[1], 0.5f - pos[2]};
    float inv = 1.0f / sqrtf(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]); for (int k = 0; k < 3; ++k) f[k] *= inv;
    float up[3] = {0, 1, 0}; float d = f[1];
    for (int k = 0; k < 3; ++k) up[k] -= f[k] * d;
    inv = 1.0f / sqrtf(up[0] * up[0] + up[1] * up[1] + up[2] * up[2]); for (int k = 0; k < 3; ++k) up[k] *= inv;
    float dn[3] = {-up[0], -up[1], -up[2]};
    float rt[3] = {dn[1] * f[2] - dn[2] * f[1], dn[2] * f[0] - dn[0] * f[2], dn[0] * f[1] - dn[1] * f[0]};
    for (int k = 0; k < 3; ++k) { x[k] = rt[k]; x[3 + k] = dn[k]; x[6 + k] = f[k]; x[9 + k] = pos[k]; }
}
// RGBA8 straight-alpha sRGB view [h][w]; step = marching step
void ref_synth_render_view(uint32_t view, uint32_t n_views, float radius, float focal, int w, int h, float step, uint32_t* out) {
    float xf[12]; ref_synth_camera(view, n_views, radius, xf);
    #pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
        float dc[3] = {((float)x + 0.5f - 0.5f * (float)w) / focal, ((float)y + 0.5f - 0.5f * (float)h) / focal, 1.0f};
        float d[3]; for (int k = 0; k < 3; ++k) d[k] = xf[k] * dc[0] + xf[3 + k] * dc[1] + xf[6 + k] * dc[2];
        float inv = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); for (int k = 0; k < 3; ++k) d[k] *= inv;
        const float* o = xf + 9;
        float t0 = 0.f, t1 = 1e30f;
        for (int k = 0; k < 3; ++k) { float iv = 1.0f / d[k], a = (0.f - o[k]) * iv, b = (1.f - o[k]) * iv; if (a > b) std::swap(a, b); t0 = fmaxf(t0, a); t1 = fminf(t1, b); }
        float T = 1.f, C[3] = {0, 0, 0};
        for (float t = t0 + 0.5f * step; t < t1 && T > 1e-4f; t += step) {
            float p[3] = {o[0] + t * d[0], o[1] + t * d[1], o[2] + t * d[2]}, sigma, rgb[3];
            field(p, d, &sigma, rgb);
            if (sigma <= 0.f) continue;
            float alpha = 1.f - expf(-sigma * step);
            for (int c = 0; c < 3; ++c) C[c] += T * alpha * rgb[c];
            T *= 1.f - alpha;
        }
        float A = 1.f - T; uint32_t px = 0;
        for (int c = 0; c < 3; ++c) {
            float l = A > 1e-6f ? C[c] / A : 0.f, s = l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f;
            px |= (uint32_t)std::min(std::max((int)(clampf(s, 0.f, 1.f) * 255.f + 0.5f), 0), 255) << (8 * c);
        }
        px |= (uint32_t)std::min(std::max((int)(A * 255.f + 0.5f), 0), 255) << 24;
        if (px == 0x00FF00FFu) px = 0x00FE00FFu;
        out[(size_t)y * w + x] = px;
    }
}
}
