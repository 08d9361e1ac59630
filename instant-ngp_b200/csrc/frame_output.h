// Output stage of a rendered frame, headless: sample accumulation, background + exposure + tonemap curve + output colour space,
// and the ground-truth overlay (reference: src/render_buffer.cu:232-346 accumulate_kernel / tonemap, :348-418 overlay_image_kernel,
// :533-565 tonemap_kernel; CudaRenderBuffer::accumulate / tonemap / overlay_image :620-700).  The GL surface and the DLSS branch of
// the reference are unavailable on this box: the "surface" here is a plain float4 device buffer.
#pragma once
#include "ngp_common.cuh"
namespace ngp {

enum ColorSpace : int { CS_LINEAR = 0, CS_SRGB = 1, CS_VISPOSNEG = 2 };                          // EColorSpace (common.h:159-163)
enum TonemapCurve : int { TM_IDENTITY = 0, TM_ACES = 1, TM_HABLE = 2, TM_REINHARD = 3 };         // ETonemapCurve (common.h:166-171)

// accumulate[i] = running mean over the samples seen so far (sample_count of them) and this frame, in `color_space`
void frame_accumulate(cudaStream_t s, uint32_t n_pixels, const float* frame_rgba, float* accumulate_rgba, float sample_count, int color_space);
// surface[i] = tonemap(accumulate[i] over background) ; background_color is sRGB RGBA
void frame_tonemap(cudaStream_t s, uint32_t n_pixels, float exposure, const float* background_rgba, const float* accumulate_rgba, int color_space,
                   int output_color_space, int tonemap_curve, bool clamp_output_color, bool unmultiply_alpha, float* surface_rgba);
// blend a training image (device pixels, RGBA8 sRGB or float linear premultiplied) over rows [y0, y1) of a w x h surface
void frame_overlay_image(cudaStream_t s, int w, int h, int y0, int y1, float alpha, const float* exposure3, const float* background_rgba,
                         int output_color_space, const void* image, int image_data_type, const int* image_res, int tonemap_curve, int color_space,
                         int fov_axis, float zoom, const float* screen_center, float* surface_rgba);

}  // namespace ngp
