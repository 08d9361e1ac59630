// Output stage of a rendered frame (see frame_output.h).  One thread per pixel, float4 loads / stores; every kernel is a single pass
// over 16 or 32 bytes per pixel and so HBM-bound by construction (a 3840x2160 frame is 133 MB of traffic per pass: ~20 us).
#include "frame_output.h"
#include "nerf_device.cuh"

namespace ngp {
namespace {

struct F4 { float x, y, z, w; };

// tonemap(x, curve) (render_buffer.cu:262-316): rational-polynomial curves with the exposure bias folded into the coefficients
__device__ void tonemap_curve3(float* c, int curve) {
    if (curve == TM_IDENTITY) return;
    for (int k = 0; k < 3; ++k) c[k] = fmaxf(c[k], 0.0f);
    float k0, k1, k2, k3, k4, k5;
    if (curve == TM_ACES) {
        k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f; k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
    } else if (curve == TM_HABLE) {
        const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f, W = 11.2f;
        k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f; k3 = A * F; k4 = B * F; k5 = D * F * F;
        const float nom = k0 * (W * W) + k1 * W + k2, denom = k3 * (W * W) + k4 * W + k5;
        const float white_scale = denom / nom;
        k0 = 4.0f * k0 * white_scale; k1 = 2.0f * k1 * white_scale; k2 = k2 * white_scale; k3 = 4.0f * k3; k4 = 2.0f * k4;
    } else {   // Reinhard on the luminance
        const float Y = 0.2126f * c[0] + 0.7152f * c[1] + 0.0722f * c[2];
        const float r = 1.f / (Y + 1.0f);
        for (int k = 0; k < 3; ++k) c[k] *= r;
        return;
    }
    for (int k = 0; k < 3; ++k) {
        const float sq = c[k] * c[k];
        c[k] = (sq * k0 + k1 * c[k] + k2) / (k3 * sq + k4 * c[k] + k5);
    }
}

// tonemap(col, exposure, curve, color_space, output_color_space) (render_buffer.cu:318-339)
__device__ void tonemap3(float* c, const float* exposure3, int curve, int color_space, int output_color_space) {
    if (color_space == CS_SRGB) for (int k = 0; k < 3; ++k) c[k] = srgb_to_linear(c[k]);
    for (int k = 0; k < 3; ++k) c[k] *= powf(2.0f, exposure3[k]);
    tonemap_curve3(c, curve);
    if (output_color_space == CS_SRGB) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
}

__global__ void accumulate_kernel(uint32_t n, const float4* __restrict__ frame, float4* __restrict__ accumulate, float sample_count, int color_space) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 color = frame[i], tmp = accumulate[i];
    const float inv = sample_count + 1;
    if (color_space == CS_VISPOSNEG) {
        const float val = color.x - color.y;
        float tmp_val = tmp.x - tmp.y;
        tmp_val = (tmp_val * sample_count + val) / inv;
        tmp.x = fmaxf(tmp_val, 0.0f); tmp.y = fmaxf(-tmp_val, 0.0f);
    } else {
        if (color_space == CS_SRGB) { color.x = linear_to_srgb(color.x); color.y = linear_to_srgb(color.y); color.z = linear_to_srgb(color.z); }
        tmp.x = (tmp.x * sample_count + color.x) / inv; tmp.y = (tmp.y * sample_count + color.y) / inv; tmp.z = (tmp.z * sample_count + color.z) / inv;
    }
    tmp.w = (tmp.w * sample_count + color.w) / inv;
    accumulate[i] = tmp;
}

__global__ void tonemap_kernel(uint32_t n, float exposure, F4 bg, const float4* __restrict__ accumulate, int color_space, int output_color_space, int curve,
                               bool clamp_output_color, bool unmultiply_alpha, float4* __restrict__ surface) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // the background colour is given in sRGB: to linear unless that is the space the frame was accumulated in
    if (color_space != CS_SRGB) { bg.x = srgb_to_linear(bg.x); bg.y = srgb_to_linear(bg.y); bg.z = srgb_to_linear(bg.z); }
    const float4 a = accumulate[i];
    const float weight = (1 - a.w) * bg.w;
    float c[3] = {a.x + bg.x * weight, a.y + bg.y * weight, a.z + bg.z * weight};
    float alpha = a.w + weight;
    const float e3[3] = {exposure, exposure, exposure};
    tonemap3(c, e3, curve, color_space, output_color_space);
    if (unmultiply_alpha && alpha > 0.0f) { c[0] /= alpha; c[1] /= alpha; c[2] /= alpha; }
    if (clamp_output_color) { for (int k = 0; k < 3; ++k) c[k] = fminf(fmaxf(c[k], 0.0f), 1.0f); alpha = fminf(fmaxf(alpha, 0.0f), 1.0f); }
    surface[i] = make_float4(c[0], c[1], c[2], alpha);
}

__global__ void overlay_image_kernel(int W, int H, int y0, int y1, float alpha, F4 exposure, F4 bg, const void* __restrict__ image, int data_type, int iw, int ih,
                                     int curve, int color_space, int output_color_space, int fov_axis, float zoom, float scx, float scy, float4* __restrict__ surface) {
    const int x = threadIdx.x + blockDim.x * blockIdx.x, y = y0 + (int)(threadIdx.y + blockDim.y * blockIdx.y);
    if (x >= W || y >= y1) return;
    const float scale = (float)(fov_axis == 0 ? iw : ih) / (float)(fov_axis == 0 ? W : H);
    float fx = x + 0.5f, fy = y + 0.5f;
    fx -= W * 0.5f; fx /= zoom; fx += scx * W;
    fy -= H * 0.5f; fy /= zoom; fy += scy * H;
    const float u = (fx - W * 0.5f) * scale + iw * 0.5f, v = (fy - H * 0.5f) * scale + ih * 0.5f;
    const int srcx = (int)floorf(u), srcy = (int)floorf(v);
    float c[4] = {0, 0, 0, 0};
    if (!(srcx >= iw || srcy >= ih || srcx < 0 || srcy < 0)) {
        const size_t idx = (size_t)srcx + (size_t)srcy * iw;
        if (data_type == IMG_BYTE) {
            const uint32_t val = __ldg(reinterpret_cast<const uint32_t*>(image) + idx);
            if (val == 0x00FF00FFu) { c[0] = c[1] = c[2] = c[3] = -1.0f; }
            else {
                const float al = (float)(val >> 24) * (1.0f / 255.0f);
                c[0] = srgb_to_linear((float)(val & 0xFF) * (1.0f / 255.0f)) * al; c[1] = srgb_to_linear((float)((val >> 8) & 0xFF) * (1.0f / 255.0f)) * al;
                c[2] = srgb_to_linear((float)((val >> 16) & 0xFF) * (1.0f / 255.0f)) * al; c[3] = al;
            }
        } else {
            const float4 p = __ldg(reinterpret_cast<const float4*>(image) + idx);
            c[0] = p.x; c[1] = p.y; c[2] = p.z; c[3] = p.w;
        }
    }
    if (color_space != CS_SRGB) { bg.x = srgb_to_linear(bg.x); bg.y = srgb_to_linear(bg.y); bg.z = srgb_to_linear(bg.z); }
    else if (c[3] > 0) { for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k] / c[3]) * c[3]; }
    else { c[0] = c[1] = c[2] = 0.0f; }
    const float weight = (1 - c[3]) * bg.w;
    c[0] += bg.x * weight; c[1] += bg.y * weight; c[2] += bg.z * weight; c[3] += weight;
    const float e3[3] = {exposure.x, exposure.y, exposure.z};
    tonemap3(c, e3, curve, color_space, output_color_space);
    const size_t o = (size_t)x + (size_t)(y - y0) * W;
    const float4 prev = surface[o];
    surface[o] = make_float4(c[0] * alpha + prev.x * (1.f - alpha), c[1] * alpha + prev.y * (1.f - alpha), c[2] * alpha + prev.z * (1.f - alpha),
                             c[3] * alpha + prev.w * (1.f - alpha));
}

}  // namespace

void frame_accumulate(cudaStream_t s, uint32_t n_pixels, const float* frame_rgba, float* accumulate_rgba, float sample_count, int color_space) {
    if (!n_pixels) return;
    accumulate_kernel<<<div_round_up(n_pixels, 256), 256, 0, s>>>(n_pixels, reinterpret_cast<const float4*>(frame_rgba), reinterpret_cast<float4*>(accumulate_rgba),
                                                                 sample_count, color_space);
    NGP_POST_LAUNCH();
}

void frame_tonemap(cudaStream_t s, uint32_t n_pixels, float exposure, const float* bg, const float* accumulate_rgba, int color_space, int output_color_space,
                   int tonemap_curve, bool clamp_output_color, bool unmultiply_alpha, float* surface_rgba) {
    if (!n_pixels) return;
    tonemap_kernel<<<div_round_up(n_pixels, 256), 256, 0, s>>>(n_pixels, exposure, F4{bg[0], bg[1], bg[2], bg[3]}, reinterpret_cast<const float4*>(accumulate_rgba),
                                                              color_space, output_color_space, tonemap_curve, clamp_output_color, unmultiply_alpha,
                                                              reinterpret_cast<float4*>(surface_rgba));
    NGP_POST_LAUNCH();
}

void frame_overlay_image(cudaStream_t s, int w, int h, int y0, int y1, float alpha, const float* e3, const float* bg, int output_color_space, const void* image,
                         int image_data_type, const int* image_res, int tonemap_curve, int color_space, int fov_axis, float zoom, const float* screen_center,
                         float* surface_rgba) {
    const dim3 threads = {16, 8, 1};
    const dim3 blocks = {div_round_up((uint32_t)w, 16u), div_round_up((uint32_t)(y1 - y0), 8u), 1};
    overlay_image_kernel<<<blocks, threads, 0, s>>>(w, h, y0, y1, alpha, F4{e3[0], e3[1], e3[2], 0.f}, F4{bg[0], bg[1], bg[2], bg[3]}, image, image_data_type,
                                                   image_res[0], image_res[1], tonemap_curve, color_space, output_color_space, fov_axis, zoom, screen_center[0],
                                                   screen_center[1], reinterpret_cast<float4*>(surface_rgba));
    NGP_POST_LAUNCH();
}

}  // namespace ngp
