// Host side of load_nerf (see nerf_dataset.h).  Control flow and arithmetic follow src/nerf_loader.cu:161-726 of the reference.
#include "nerf_dataset.h"
#include "ngp_common.cuh"
#include "nerf_device.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>
#include <sys/stat.h>

namespace ngp {

static bool file_exists(const std::string& p) { struct stat st; return !p.empty() && stat(p.c_str(), &st) == 0; }
static std::string parent_path(const std::string& p) { const size_t k = p.find_last_of('/'); return k == std::string::npos ? std::string(".") : (k == 0 ? std::string("/") : p.substr(0, k)); }
static bool has_extension(const std::string& p) {
    const size_t slash = p.find_last_of('/'), dot = p.find_last_of('.');
    return dot != std::string::npos && (slash == std::string::npos || dot > slash + 1) && dot + 1 < p.size();
}

bool natural_less(const std::string& a, const std::string& b) {
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size()) {
        if (std::isdigit((unsigned char)a[i]) && std::isdigit((unsigned char)b[j])) {
            size_t i0 = i, j0 = j;
            while (i0 < a.size() && a[i0] == '0') ++i0;
            while (j0 < b.size() && b[j0] == '0') ++j0;
            size_t i1 = i0, j1 = j0;
            while (i1 < a.size() && std::isdigit((unsigned char)a[i1])) ++i1;
            while (j1 < b.size() && std::isdigit((unsigned char)b[j1])) ++j1;
            if (i1 - i0 != j1 - j0) return i1 - i0 < j1 - j0;            // more significant digits -> larger
            const int c = a.compare(i0, i1 - i0, b, j0, j1 - j0);
            if (c != 0) return c < 0;
            i = i1; j = j1;
        } else {
            if (a[i] != b[j]) return (unsigned char)a[i] < (unsigned char)b[j];
            ++i; ++j;
        }
    }
    return a.size() - i < b.size() - j;
}

static const char* SUPPORTED_FORMATS[] = {"png", "jpg", "jpeg", "bmp", "gif", "tga", "pic", "pnm", "psd", "exr"};   // nerf_loader.cu:302-304

static std::string resolve_path(const std::string& base, const std::string& local) {   // nerf_loader.cu:306-317
    std::string path = (!local.empty() && local[0] == '/') ? local : base + "/" + local;
    if (!has_extension(path) && !file_exists(path))
        for (const char* f : SUPPORTED_FORMATS) if (file_exists(path + "." + f)) return path + "." + f;
    return path;
}

static float num(const Json& j) {
    if (j.type != Json::Number) throw Error("transforms.json: expected a number");
    return (float)j.num;
}
static float fov_to_focal_length(int resolution, float degrees) { return 0.5f * (float)resolution / tanf(0.5f * degrees * 3.14159265358979323846f / 180.0f); }

struct LensState { int mode = LENS_PERSPECTIVE; float params[4] = {0, 0, 0, 0}; float principal_point[2] = {0.5f, 0.5f}; float rolling_shutter[4] = {0, 0, 0, 0}; };

// read_lens (nerf_loader.cu:161-227), the modes this path implements (Perspective, OpenCV)
static void read_lens(const Json& json, LensState& l) {
    int mode = LENS_PERSPECTIVE;
    const bool fisheye = json.value("is_fisheye", false);
    auto read = [&](const char* name, int idx) {
        if (json.contains(name)) { l.params[idx] = num(json.at(name)); if (l.params[idx] != 0.f) mode = fisheye ? -1 : LENS_OPENCV; }
    };
    read("k1", 0); read("k2", 1); read("k3", 2); read("k4", 3); read("p1", 2); read("p2", 3);
    if (json.contains("cx")) l.principal_point[0] = num(json.at("cx")) / num(json.at("w"));
    if (json.contains("cy")) l.principal_point[1] = num(json.at("cy")) / num(json.at("h"));
    if (json.contains("rolling_shutter")) {   // [A, B, C(, D)]: pixel time = A + B u + C v + D motionblur_time (:190-202)
        const Json& rs = json.at("rolling_shutter");
        const float d = rs.arr.size() >= 4 ? num(rs.arr[3]) : 0.f;
        l.rolling_shutter[0] = num(rs.arr.at(0)); l.rolling_shutter[1] = num(rs.arr.at(1)); l.rolling_shutter[2] = num(rs.arr.at(2)); l.rolling_shutter[3] = d;
    }
    if (json.contains("ftheta_p0") || json.contains("latlong") || json.contains("equirectangular") || mode == -1)
        throw Error("transforms.json: only the Perspective and OpenCV lens models are implemented on this path");
    if (mode != LENS_PERSPECTIVE) l.mode = mode;
}

// read_focal_length (nerf_loader.cu:229-258)
static bool read_focal_length(const Json& json, float* focal, const int* res) {
    auto one = [&](int resolution, const std::string& axis) -> float {
        if (json.contains(axis + "_fov")) return fov_to_focal_length(resolution, num(json.at(axis + "_fov")));
        if (json.contains("fl_" + axis)) return num(json.at("fl_" + axis));
        if (json.contains("camera_angle_" + axis)) return fov_to_focal_length(resolution, num(json.at("camera_angle_" + axis)) * 180 / 3.14159265358979323846f);
        return 0.0f;
    };
    const float x = one(res[0], "x"), y = one(res[1], "y");
    if (x != 0) { focal[0] = focal[1] = x; if (y != 0) focal[1] = y; }
    else if (y != 0) focal[0] = focal[1] = y;
    else return false;
    return true;
}

// nerf_matrix_to_ngp (nerf_loader.h:120-140): columns 1, 2 negated, position scaled and offset, rows cycled xyz <- yzx
void NerfDataset::nerf_matrix_to_ngp(const float* nerf, float* ngp) const {
    float m[12]; memcpy(m, nerf, sizeof(m));
    for (int r = 0; r < 3; ++r) { m[3 + r] *= -1.f; m[6 + r] *= -1.f; m[9 + r] = m[9 + r] * scale + offset[r]; }
    if (from_mitsuba) { for (int r = 0; r < 3; ++r) { m[r] *= -1.f; m[6 + r] *= -1.f; } memcpy(ngp, m, sizeof(m)); return; }
    for (int c = 0; c < 4; ++c) { ngp[c * 3 + 0] = m[c * 3 + 1]; ngp[c * 3 + 1] = m[c * 3 + 2]; ngp[c * 3 + 2] = m[c * 3 + 0]; }
}

NerfDataset::NerfDataset(const std::vector<std::string>& json_paths) {
    if (json_paths.empty()) throw Error("Cannot load NeRF data from an empty set of paths.");
    for (const std::string& p : json_paths) {
        std::ifstream f(p);
        if (!f) throw Error("NeRF dataset '" + p + "' does not exist.");
        std::stringstream ss; ss << f.rdbuf();
        m_jsons.push_back(Json::parse(ss.str()));
    }
    // pass 1 (nerf_loader.cu:319-380): frame order and culling per json
    std::vector<std::vector<Json>> kept(m_jsons.size());
    for (size_t ji = 0; ji < m_jsons.size(); ++ji) {
        Json& json = m_jsons[ji];
        const std::string base = parent_path(json_paths[ji]);
        if (!json.contains("frames") || json.at("frames").type != Json::Array) continue;
        std::vector<Json> frames = json.at("frames").arr;
        for (Json& fr : frames) {
            if (!fr.contains("file_path") || fr.at("file_path").type != Json::String) throw Error("transforms.json: frame without file_path");
            std::replace(fr["file_path"].str.begin(), fr["file_path"].str.end(), '\\', '/');
        }
        const float sharpness_discard_threshold = (float)json.value("sharpness_discard_threshold", 0.0);
        std::stable_sort(frames.begin(), frames.end(), [](const Json& a, const Json& b) { return natural_less(a.at("file_path").str, b.at("file_path").str); });
        if (json.contains("n_frames")) frames.resize(std::min(frames.size(), (size_t)json.at("n_frames").num));
        if (!frames.empty() && frames[0].contains("sharpness")) {
            std::vector<Json> copy; copy.swap(frames);
            const int neighborhood_size = 3, n = (int)copy.size();
            for (int i = 0; i < n; ++i) {   // "Kill blurrier frames than their neighbors"
                float mean_sharpness = 0.0f;
                const int mean_start = std::max(0, i - neighborhood_size), mean_end = std::min(i + neighborhood_size, n - 1);
                for (int j = mean_start; j < mean_end; ++j) mean_sharpness += (float)copy[j].value("sharpness", 1.0);
                mean_sharpness /= (float)(mean_end - mean_start);
                if (file_exists(resolve_path(base, copy[i].at("file_path").str)) && copy[i].value("sharpness", 1.0) > (double)(sharpness_discard_threshold * mean_sharpness))
                    frames.push_back(copy[i]);
            }
        }
        kept[ji] = frames;
    }
    size_t total = 0; for (auto& k : kept) total += k.size();
    if (total == 0) throw Error("No training images were found for NeRF training!");
    bool enable_depth_loading = true; float depth_scale = -1.f;   // persist from one json to the next, like the transparency flags (:283-285,404)
    // pass 2 (nerf_loader.cu:405-700): scene-level fields, then per-frame lens / pose
    for (size_t ji = 0; ji < m_jsons.size(); ++ji) {
        const Json& json = m_jsons[ji];
        const std::string base = parent_path(json_paths[ji]);
        if (json.contains("normal_mts_args")) from_mitsuba = true;
        if (from_mitsuba) { scale = 0.66f; offset[0] = offset[1] = offset[2] = 0.25f * scale; }
        if (json.contains("render_aabb")) {
            has_render_aabb = true;
            for (int k = 0; k < 3; ++k) { render_aabb[k] = num(json.at("render_aabb").arr.at(0).arr.at(k)); render_aabb[3 + k] = num(json.at("render_aabb").arr.at(1).arr.at(k)); }
        }
        if (json.contains("white_transparent")) white_transparent = json.value("white_transparent", false);
        if (json.contains("black_transparent")) black_transparent = json.value("black_transparent", false);
        if (json.contains("scale")) scale = num(json.at("scale"));
        if (json.contains("enable_depth_loading")) enable_depth_loading = json.value("enable_depth_loading", true);   // :420-423
        if (json.contains("integer_depth_scale")) depth_scale = num(json.at("integer_depth_scale"));                  // :471-473
        LensState lens;
        read_lens(json, lens);
        if (json.contains("aabb_scale")) aabb_scale = (int)num(json.at("aabb_scale"));
        // per-view latent codes (nerf_loader.cu:512-514) widen the network's direction input: not part of this library, refused rather than ignored
        NGP_CHECK(!json.contains("n_extra_learnable_dims") || num(json.at("n_extra_learnable_dims")) <= 0.f, "n_extra_learnable_dims > 0 (per-view latent codes) is not supported");
        if (json.contains("offset")) {
            const Json& o = json.at("offset");
            for (int k = 0; k < 3; ++k) offset[k] = o.type == Json::Array ? num(o.arr.at(k)) : num(o);
        }
        if (json.contains("aabb")) {   // isotropic fit of the given box into the unit cube (nerf_loader.cu:489-495)
            const Json& a = json.at("aabb");
            float lo[3], hi[3]; for (int k = 0; k < 3; ++k) { lo[k] = num(a.arr.at(0).arr.at(k)); hi[k] = num(a.arr.at(1).arr.at(k)); }
            const float length = std::max(0.000001f, std::max(std::max(std::fabs(hi[0] - lo[0]), std::fabs(hi[1] - lo[1])), std::fabs(hi[2] - lo[2])));
            scale = 1.f / length;
            for (int k = 0; k < 3; ++k) offset[k] = ((hi[k] + lo[k]) * 0.5f) * -scale + 0.5f;
        }
        if (json.contains("up")) { up[0] = num(json.at("up").arr.at(1)); up[1] = num(json.at("up").arr.at(2)); up[2] = num(json.at("up").arr.at(0)); }
        for (size_t fi = 0; fi < kept[ji].size(); ++fi) {
            const Json& fr = kept[ji][fi];
            DatasetFrame d;
            d.json_index = ji; d.frame_index = fi; d.white_transparent = white_transparent; d.black_transparent = black_transparent;
            d.file_path = fr.at("file_path").str;
            d.resolved_path = resolve_path(base, d.file_path);
            d.depth_scale = depth_scale;
            if (enable_depth_loading && depth_scale > 0.f && fr.contains("depth_path")) {   // :609-621
                const std::string dp = resolve_path(base, fr.at("depth_path").str);
                if (file_exists(dp)) d.depth_path = dp;
            }
            const Json& mat = fr.contains("transform_matrix_start") ? fr.at("transform_matrix_start") : fr.at("transform_matrix");
            const Json& mat_end = fr.contains("transform_matrix_end") ? fr.at("transform_matrix_end") : mat;   // :648-649
            for (int m = 0; m < 3; ++m) for (int n = 0; n < 4; ++n) { d.xform_nerf[n * 3 + m] = num(mat.arr.at(m).arr.at(n)); d.xform_end_nerf[n * 3 + m] = num(mat_end.arr.at(m).arr.at(n)); }
            LensState l = lens;           // "set these from the base settings", then "see if there is a per-frame override"
            read_lens(fr, l);
            d.lens_mode = l.mode; memcpy(d.lens_params, l.params, sizeof(l.params)); memcpy(d.principal_point, l.principal_point, sizeof(l.principal_point));
            memcpy(d.rolling_shutter, l.rolling_shutter, sizeof(l.rolling_shutter));
            nerf_matrix_to_ngp(d.xform_nerf, d.xform);
            frames.push_back(std::move(d));
        }
    }
    // scale / offset of a later json apply to every frame in the reference too (conversion runs after all jsons are read)
    for (DatasetFrame& d : frames) { nerf_matrix_to_ngp(d.xform_nerf, d.xform); nerf_matrix_to_ngp(d.xform_end_nerf, d.xform_end); }
}

void NerfDataset::set_depth(size_t i, const uint16_t* pixels, int w, int h) {
    NGP_CHECK(i < frames.size(), "Invalid frame index");
    DatasetFrame& d = frames[i];
    NGP_CHECK(pixels, "Could not load depth image '" + d.depth_path + "'.");
    if (w != d.res[0] || h != d.res[1]) throw Error("Depth image " + d.depth_path + " has wrong resolution.");
    d.depth_pixels.assign(pixels, pixels + (size_t)w * h);
}

void NerfDataset::set_image(size_t i, const void* pixels, int data_type, int w, int h) {
    NGP_CHECK(i < frames.size(), "Invalid frame index");
    NGP_CHECK(pixels && w > 0 && h > 0, "set_image: bad image");
    NGP_CHECK(data_type == IMG_BYTE || data_type == IMG_FLOAT, "set_image: data type must be RGBA8 (0) or RGBA float32 (1)");
    DatasetFrame& d = frames[i];
    d.res[0] = w; d.res[1] = h; d.data_type = data_type;
    const size_t n = (size_t)w * h, bytes = n * (data_type == IMG_BYTE ? 4 : 16);
    d.pixels.assign((const uint8_t*)pixels, (const uint8_t*)pixels + bytes);
    if (data_type == IMG_BYTE && (d.white_transparent || d.black_transparent)) {   // convert_rgba32 (nerf_loader.cu:49-71)
        for (size_t p = 0; p < n; ++p) {
            uint8_t* q = &d.pixels[p * 4];
            if (d.white_transparent && q[0] == 255 && q[1] == 255 && q[2] == 255) q[3] = 0;
            if (d.black_transparent && q[0] == 0 && q[1] == 0 && q[2] == 0) q[3] = 0;
        }
    }
    const Json& json = m_jsons[d.json_index];
    bool got = read_focal_length(json, d.focal_length, d.res);
    // the frame as kept after sorting / culling is looked up by path (the json's own array is left in file order here)
    for (const Json& fr : json.at("frames").arr) {
        if (!fr.contains("file_path")) continue;
        std::string p = fr.at("file_path").str; std::replace(p.begin(), p.end(), '\\', '/');
        if (p == d.file_path) { got |= read_focal_length(fr, d.focal_length, d.res); break; }
    }
    if (!got) throw Error("Couldn't read fov.");
    d.has_focal = true;
}

bool NerfDataset::complete() const {
    for (const DatasetFrame& d : frames) if (!d.has_focal || d.pixels.empty()) return false;
    return !frames.empty();
}

}  // namespace ngp
