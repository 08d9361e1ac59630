// Procedural analytic NeRF scene (BASELINE.json: "100 posed 800x800 views of a procedural analytic density field"):
// three soft spheres + a slab with view-dependent colour inside the unit cube, ray-marched at sqrt(3)/4096 steps.
// Shared by the CUDA data synthesiser (synth_scene.cu) and nothing else; the CPU oracle carries its own copy.
#pragma once
#include "ngp_common.cuh"

namespace ngp {

struct SynthPrim { float c[3]; float r[3]; float rgb[3]; int box; };

NGP_HD float synth_smooth(float x) { x = clampf(x, 0.f, 1.f); return x * x * (3.f - 2.f * x); }

NGP_HD void synth_field(V3 p, V3 dir, float& sigma, V3& rgb) {
    const SynthPrim prims[4] = {
        {{0.50f, 0.47f, 0.50f}, {0.17f, 0.17f, 0.17f}, {0.90f, 0.30f, 0.20f}, 0},
        {{0.68f, 0.62f, 0.42f}, {0.10f, 0.10f, 0.10f}, {0.20f, 0.80f, 0.30f}, 0},
        {{0.34f, 0.64f, 0.60f}, {0.08f, 0.08f, 0.08f}, {0.20f, 0.40f, 0.90f}, 0},
        {{0.50f, 0.26f, 0.50f}, {0.27f, 0.03f, 0.27f}, {0.85f, 0.80f, 0.65f}, 1},
    };
    sigma = 0.f; rgb = v3(0, 0, 0);
    for (int k = 0; k < 4; ++k) {
        const SynthPrim& q = prims[k];
        float sd;   // signed distance-like: negative inside
        if (q.box) {
            const float dx = fabsf(p.x - q.c[0]) - q.r[0], dy = fabsf(p.y - q.c[1]) - q.r[1], dz = fabsf(p.z - q.c[2]) - q.r[2];
            sd = fmaxf(fmaxf(dx, dy), dz);
        } else {
            const float dx = p.x - q.c[0], dy = p.y - q.c[1], dz = p.z - q.c[2];
            sd = sqrtf(dx * dx + dy * dy + dz * dz) - q.r[0];
        }
        const float s = 400.f * synth_smooth(0.5f - sd / 0.02f);
        if (s > 0.f) {
            // view dependence: a low-order lobe in the viewing direction (SH-2 like)
            const float vd = 0.8f + 0.2f * (dir.x * 0.3f + dir.y * 0.8f + dir.z * 0.5f) * (dir.x * 0.3f + dir.y * 0.8f + dir.z * 0.5f);
            rgb = rgb + s * vd * v3(q.rgb[0], q.rgb[1], q.rgb[2]);
            sigma += s;
        }
    }
    if (sigma > 0.f) rgb = rgb * (1.0f / sigma);
}

// camera i of n on an upper-hemisphere Fibonacci lattice of radius `radius` around (0.5,0.5,0.5), +y up, looking at the centre.
// xform: mat4x3 column major (right, down, forward, position), right x down = forward.
NGP_HD void synth_camera(uint32_t i, uint32_t n, float radius, float* xform) {
    const float golden = 2.39996322972865332f;
    const float h = 0.08f + 0.87f * ((float)i + 0.5f) / (float)n;   // height fraction in (0.08, 0.95)
    const float r = sqrtf(fmaxf(1.f - h * h, 0.f));
    const float phi = golden * (float)i;
    const V3 pos = v3(0.5f + radius * r * cosf(phi), 0.5f + radius * h, 0.5f + radius * r * sinf(phi));
    const V3 fwd = normalize3(v3(0.5f, 0.5f, 0.5f) - pos);
    V3 up = v3(0, 1, 0);
    up = normalize3(up - fwd * dot3(up, fwd));
    const V3 down = v3(-up.x, -up.y, -up.z);
    const V3 right = v3(down.y * fwd.z - down.z * fwd.y, down.z * fwd.x - down.x * fwd.z, down.x * fwd.y - down.y * fwd.x);
    xform[0] = right.x; xform[1] = right.y; xform[2] = right.z; xform[3] = down.x; xform[4] = down.y; xform[5] = down.z;
    xform[6] = fwd.x; xform[7] = fwd.y; xform[8] = fwd.z; xform[9] = pos.x; xform[10] = pos.y; xform[11] = pos.z;
}

// one pixel of the ground-truth view -> RGBA8 (straight-alpha sRGB, the loader's Byte convention: common_device.cuh:885-913)
NGP_HD uint32_t synth_pixel(const float* xform, float focal, int w, int h, int x, int y, float step) {
    const V3 dcam = v3(((float)x + 0.5f - 0.5f * (float)w) / focal, ((float)y + 0.5f - 0.5f * (float)h) / focal, 1.0f);
    const V3 d = normalize3(mat3_mul(xform, dcam));
    const V3 o = v3(xform[9], xform[10], xform[11]);
    float t0 = 0.f, t1 = 1e30f;
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    for (int k = 0; k < 3; ++k) {
        const float inv = 1.0f / dd[k];
        float a = (0.f - oo[k]) * inv, b = (1.f - oo[k]) * inv;
        if (a > b) { const float tmp = a; a = b; b = tmp; }
        t0 = fmaxf(t0, a); t1 = fminf(t1, b);
    }
    float T = 1.f; V3 C = v3(0, 0, 0);
    for (float t = t0 + 0.5f * step; t < t1 && T > 1e-4f; t += step) {
        float sigma; V3 rgb; synth_field(ray_at(o, d, t), d, sigma, rgb);
        if (sigma <= 0.f) continue;
        const float alpha = 1.f - expf(-sigma * step);
        C = C + (T * alpha) * rgb; T *= 1.f - alpha;
    }
    const float A = 1.f - T;
    V3 straight = A > 1e-6f ? C * (1.0f / A) : v3(0, 0, 0);
    auto enc = [](float l) { float s = l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; return (uint32_t)clampi((int)(clampf(s, 0.f, 1.f) * 255.f + 0.5f), 0, 255); };
    uint32_t a8 = (uint32_t)clampi((int)(A * 255.f + 0.5f), 0, 255);
    uint32_t px = enc(straight.x) | (enc(straight.y) << 8) | (enc(straight.z) << 16) | (a8 << 24);
    if (px == 0x00FF00FFu) px = 0x00FE00FFu;   // never emit the loader's "masked pixel" sentinel
    return px;
}

}  // namespace ngp
