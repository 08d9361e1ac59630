// NeRF dataset description: the host side of load_nerf (reference: src/nerf_loader.cu:161-258 read_lens / read_focal_length,
// :259-726 load_nerf; include/neural-graphics-primitives/nerf_loader.h:110-171 NerfDataset).  Everything load_nerf derives from a
// transforms.json — frame order, sharpness culling, intrinsics, lens, NeRF -> NGP pose conversion, scale / offset / aabb_scale —
// is done here; decoding the image files (stb_image in the reference, a third-party dependency) is left to the caller, who hands
// decoded RGBA8 / RGBA float pixels to set_image().
#pragma once
#include "json_mini.h"
#include <string>
#include <vector>

namespace ngp {

struct DatasetFrame {
    std::string file_path;        // as written in the json ('\\' -> '/')
    std::string resolved_path;    // base directory applied, extension searched (nerf_loader.cu:306-317)
    float xform_nerf[12];         // transform_matrix rows 0-2 as a column-major mat4x3 (nerf_loader.cu:679-684)
    float xform[12];              // nerf_matrix_to_ngp (nerf_loader.h:120-140)
    float xform_end_nerf[12], xform_end[12];   // transform_matrix_end (default: the start pose), same conversions
    float rolling_shutter[4] = {0, 0, 0, 0};   // "rolling_shutter" of the scene or the frame (nerf_loader.cu:190-202)
    float focal_length[2] = {0, 0}; bool has_focal = false;
    float principal_point[2] = {0.5f, 0.5f};
    int lens_mode = 0; float lens_params[4] = {0, 0, 0, 0};
    int res[2] = {0, 0};
    int data_type = 0;            // 0 RGBA8, 1 RGBA float32
    std::vector<uint8_t> pixels;  // converted copy (white/black-transparent rules of convert_rgba32, nerf_loader.cu:49-71)
    size_t json_index = 0, frame_index = 0;   // which json / which frame of it (after sorting and culling)
    bool white_transparent = false, black_transparent = false;   // the scene flags as they stood when this frame's json was read (LoadedImageInfo defaults)
    // depth supervision (nerf_loader.cu:471-473,609-621): "integer_depth_scale" as it stood when this frame's json was read (-1: none), the frame's
    // resolved "depth_path" (empty: no file to read) and its 16-bit pixels once the caller has decoded them
    float depth_scale = -1.f;
    std::string depth_path;
    std::vector<uint16_t> depth_pixels;
};

class NerfDataset {
public:
    // one or more transforms.json files describing the same scene (load_nerf takes a vector of paths)
    explicit NerfDataset(const std::vector<std::string>& json_paths);
    size_t n_images() const { return frames.size(); }
    // decoded pixels of frame i: [h][w][4] uint8 (sRGB, straight alpha) or float32; focal length is resolved here because the
    // fov forms need the resolution (read_focal_length, nerf_loader.cu:229-258)
    void set_image(size_t i, const void* pixels, int data_type, int w, int h);
    // decoded depth map of frame i: [h][w] uint16, same resolution as the image (set_image first)
    void set_depth(size_t i, const uint16_t* pixels, int w, int h);
    bool complete() const;

    std::vector<DatasetFrame> frames;
    float scale = 0.33f, offset[3] = {0.5f, 0.5f, 0.5f};   // NERF_SCALE (nerf_loader.h:29), nerf_loader.cu:392-393
    int aabb_scale = 1;
    bool from_mitsuba = false, white_transparent = false, black_transparent = false, has_render_aabb = false;
    float render_aabb[6] = {0, 0, 0, 1, 1, 1};
    float up[3] = {0, 1, 0};
    void nerf_matrix_to_ngp(const float* nerf, float* ngp) const;
private:
    std::vector<Json> m_jsons;
};

// SI::natural::compare as load_nerf uses it to order frames by file_path: digit runs compare by value
bool natural_less(const std::string& a, const std::string& b);

}  // namespace ngp
