// extern "C" boundary (include/ngp_b200.h): operator-level entry points.  Exceptions never cross the ABI — they are
// caught, stored per thread and reported as a non-zero status + ngp_last_error().
#include <sys/stat.h>
#include "../../include/ngp_b200.h"
#include "testbed.h"
#include "msgpack_mini.h"
#include "nerf_dataset.h"
#include <cstring>

namespace ngp { std::atomic<uint64_t> g_launch_count{0}; }

static thread_local std::string g_last_error;

#define NGP_API_BEGIN try {
#define NGP_API_END                                            \
    return 0;                                                  \
    } catch (const std::exception& e) { g_last_error = e.what(); return 1; } \
      catch (...) { g_last_error = "unknown error"; return 1; }

using namespace ngp;

struct ngp_encoding { GridLayout g; const float* max_level = nullptr; };
struct ngp_nerf_network { NerfNetwork net; explicit ngp_nerf_network(const Json& j) : net(j) {} };
struct ngp_nerf_dataset { NerfDataset ds; explicit ngp_nerf_dataset(const std::vector<std::string>& p) : ds(p) {} };
struct ngp_encoded_mlp { EncodedMlp net; ngp_encoded_mlp(const Json& j, uint32_t i, uint32_t o) : net(j, i, o) {} };

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }
static inline const __half* H(const void* p) { return reinterpret_cast<const __half*>(p); }
static inline __half* H(void* p) { return reinterpret_cast<__half*>(p); }
static inline const float* F(const void* p) { return reinterpret_cast<const float*>(p); }
static inline float* F(void* p) { return reinterpret_cast<float*>(p); }

static void require_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) { cudaGetLastError(); throw Error("ngp_b200: no CUDA device available (this library has no CPU fallback)"); }
}

// used by noalloc.cu (the allocation-free inference entry point lives beside its static scratch array)
namespace ngp {
NerfNetwork* capi_nerf_network(ngp_nerf_network_t h) { return &h->net; }
void capi_set_last_error(const char* msg) { g_last_error = msg ? msg : "unknown error"; }
}

extern "C" {

const char* ngp_last_error(void) { return g_last_error.c_str(); }
int ngp_abi_version(void) { return NGP_B200_ABI_VERSION; }
uint64_t ngp_kernel_launch_count(void) { return g_launch_count.load(); }
int ngp_host_has_mask_pixels(const void* rgba8_pixels, uint64_t n_pixels, int32_t* has_mask) {
    NGP_API_BEGIN NGP_CHECK(has_mask && (rgba8_pixels || n_pixels == 0), "null argument");
    *has_mask = host_has_mask_pixels(rgba8_pixels, (size_t)n_pixels) ? 1 : 0;
    NGP_API_END
}
void ngp_profile_enable(int on) { profile_enable(on != 0); }
int ngp_profile_report(char* buf, uint64_t buf_size, int reset) {
    NGP_API_BEGIN NGP_CHECK(buf && buf_size > 2, "null argument");
    const std::string r = profile_report_json(reset != 0);
    NGP_CHECK(r.size() + 1 <= buf_size, "profile report buffer too small");
    memcpy(buf, r.c_str(), r.size() + 1);
    NGP_API_END
}

int ngp_profile_timeline(char* buf, uint64_t buf_size) {
    NGP_API_BEGIN NGP_CHECK(buf && buf_size > 2, "null argument");
    const std::string r = profile_timeline_json();
    NGP_CHECK(r.size() + 1 <= buf_size, "profile timeline buffer too small");
    memcpy(buf, r.c_str(), r.size() + 1);
    NGP_API_END
}
int ngp_microbench_l2_gather(void* stream, uint64_t buffer_bytes, uint32_t reps, double* out) {
    NGP_API_BEGIN NGP_CHECK(out, "null argument"); require_device();
    *out = microbench_l2_gather(S(stream), buffer_bytes, reps);
    NGP_API_END
}
int ngp_microbench_l2_red(void* stream, uint64_t buffer_bytes, int32_t vec, uint32_t reps, double* out) {
    NGP_API_BEGIN NGP_CHECK(out, "null argument"); require_device();
    *out = microbench_l2_red(S(stream), buffer_bytes, vec, reps);
    NGP_API_END
}

// ---- encoding ----
int ngp_encoding_create(const char* json, uint32_t n_dims, ngp_encoding_t* out) {
    NGP_API_BEGIN
    NGP_CHECK(json && out, "null argument");
    auto* e = new ngp_encoding();
    try { grid_from_json(e->g, Json::parse(json), n_dims); } catch (...) { delete e; throw; }
    *out = e;
    NGP_API_END
}
void ngp_encoding_destroy(ngp_encoding_t e) { delete e; }
uint64_t ngp_encoding_n_params(ngp_encoding_t e) { return e ? e->g.n_params : 0; }
uint32_t ngp_encoding_padded_output_width(ngp_encoding_t e) { return e ? e->g.padded_width : 0; }
uint32_t ngp_encoding_level_table(ngp_encoding_t e, uint32_t* offsets, float* scales, uint32_t* resolutions) {
    if (!e) return 0;
    for (uint32_t l = 0; l < e->g.n_levels; ++l) {
        if (offsets) offsets[l] = e->g.offsets[l];
        if (scales) scales[l] = e->g.scales[l];
        if (resolutions) resolutions[l] = e->g.resolutions[l];
    }
    if (offsets) offsets[e->g.n_levels] = e->g.offsets[e->g.n_levels];
    return e->g.n_levels;
}
int ngp_encoding_initialize_params(ngp_encoding_t e, void* stream, uint64_t seed, void* p) {
    NGP_API_BEGIN require_device(); NGP_CHECK(e && p, "null argument");
    grid_init(S(stream), e->g, Pcg32(seed), F(p));
    NGP_API_END
}
int ngp_encoding_forward(ngp_encoding_t e, void* stream, const void* params, const void* x, uint32_t x_stride, uint32_t n, void* y, int y_layout) {
    NGP_API_BEGIN require_device(); NGP_CHECK(e && (n == 0 || (params && x && y)), "null argument");
    grid_forward(S(stream), e->g, H(params), F(x), x_stride, n, H(y), y_layout, nullptr, e->max_level);
    NGP_API_END
}
int ngp_encoding_backward(ngp_encoding_t e, void* stream, const void* x, uint32_t x_stride, uint32_t n, const void* dy, int dy_layout, void* grad) {
    NGP_API_BEGIN require_device(); NGP_CHECK(e && (n == 0 || (x && dy && grad)), "null argument");
    grid_backward(S(stream), e->g, F(x), x_stride, n, H(dy), dy_layout, H(grad), e->max_level);
    NGP_API_END
}
int ngp_encoding_set_max_level_gpu(ngp_encoding_t e, const float* dev_max_level) {
    NGP_API_BEGIN NGP_CHECK(e, "null argument");
    e->max_level = dev_max_level;
    NGP_API_END
}
int ngp_encoding_backward_input(ngp_encoding_t e, void* stream, const void* params, const void* x, uint32_t x_stride, uint32_t n, const void* dy,
                                int dy_layout, void* dx) {
    NGP_API_BEGIN require_device(); NGP_CHECK(e && params && x && dy && dx, "null argument");
    grid_backward_input(S(stream), e->g, H(params), F(x), x_stride, n, H(dy), dy_layout, F(dx));
    NGP_API_END
}
int ngp_encoding_corners(ngp_encoding_t e, void* stream, const void* x, uint32_t x_stride, uint32_t n, uint32_t level, void* idx, void* w) {
    NGP_API_BEGIN require_device(); NGP_CHECK(e && x && idx && w, "null argument");
    grid_corners(S(stream), e->g, F(x), x_stride, n, level, reinterpret_cast<uint32_t*>(idx), F(w));
    NGP_API_END
}

// ---- NerfNetwork ----
int ngp_nerf_network_create(const char* json, ngp_nerf_network_t* out) {
    NGP_API_BEGIN NGP_CHECK(json && out, "null argument");
    *out = new ngp_nerf_network(Json::parse(json));
    NGP_API_END
}
void ngp_nerf_network_destroy(ngp_nerf_network_t n) { delete n; }
uint64_t ngp_nerf_network_n_params(ngp_nerf_network_t n) { return n ? n->net.n_params : 0; }
int ngp_nerf_network_layout(ngp_nerf_network_t n, uint64_t* od, uint64_t* orgb, uint64_t* oe, uint64_t* nm) {
    NGP_API_BEGIN NGP_CHECK(n, "null argument");
    if (od) *od = n->net.off_density; if (orgb) *orgb = n->net.off_rgb; if (oe) *oe = n->net.off_grid; if (nm) *nm = n->net.n_matrix_params;
    NGP_API_END
}
int ngp_nerf_network_initialize_params(ngp_nerf_network_t n, void* stream, uint64_t seed, void* p) {
    NGP_API_BEGIN require_device(); NGP_CHECK(n && p, "null argument");
    n->net.initialize_params(seed, F(p), S(stream));
    NGP_API_END
}
int ngp_nerf_network_inference(ngp_nerf_network_t n, void* stream, const void* params, const void* coords, uint32_t count, void* out) {
    NGP_API_BEGIN require_device(); NGP_CHECK(n && params && coords && out, "null argument");
    n->net.inference(S(stream), H(params), F(coords), count, H(out));
    NGP_API_END
}
int ngp_nerf_network_density(ngp_nerf_network_t n, void* stream, const void* params, const void* pos, uint32_t pos_stride, uint32_t count, void* out) {
    NGP_API_BEGIN require_device(); NGP_CHECK(n && params && pos && out, "null argument");
    n->net.density(S(stream), H(params), F(pos), pos_stride, count, H(out));
    NGP_API_END
}
int ngp_nerf_network_forward_backward(ngp_nerf_network_t n, void* stream, const void* params, const void* coords, uint32_t count, const void* dL,
                                      void* grads) {
    NGP_API_BEGIN require_device(); NGP_CHECK(n && params && coords && dL && grads, "null argument");
    n->net.forward_backward(S(stream), H(params), F(coords), count, H(dL), H(grads));
    NGP_API_END
}
int ngp_nerf_network_forward_backward_input(ngp_nerf_network_t n, void* stream, const void* params, const void* coords, uint32_t count, const void* dL,
                                            void* grads, void* dL_dcoords) {
    NGP_API_BEGIN require_device(); NGP_CHECK(n && params && coords && dL && grads && dL_dcoords, "null argument");
    n->net.forward_backward(S(stream), H(params), F(coords), count, H(dL), H(grads), F(dL_dcoords));
    NGP_API_END
}

// ---- NetworkWithInputEncoding ----
int ngp_encoded_mlp_create(const char* json, uint32_t n_in, uint32_t n_out, ngp_encoded_mlp_t* out) {
    NGP_API_BEGIN NGP_CHECK(json && out, "null argument");
    *out = new ngp_encoded_mlp(Json::parse(json), n_in, n_out);
    NGP_API_END
}
void ngp_encoded_mlp_destroy(ngp_encoded_mlp_t m) { delete m; }
uint64_t ngp_encoded_mlp_n_params(ngp_encoded_mlp_t m) { return m ? m->net.n_params : 0; }
int ngp_encoded_mlp_layout(ngp_encoded_mlp_t m, uint64_t* om, uint64_t* oe, uint64_t* nm) {
    NGP_API_BEGIN NGP_CHECK(m, "null argument");
    if (om) *om = m->net.off_mlp; if (oe) *oe = m->net.off_grid; if (nm) *nm = m->net.n_matrix_params;
    NGP_API_END
}
int ngp_encoded_mlp_initialize_params(ngp_encoded_mlp_t m, void* stream, uint64_t seed, void* p) {
    NGP_API_BEGIN require_device(); NGP_CHECK(m && p, "null argument");
    m->net.initialize_params(seed, F(p), S(stream));
    NGP_API_END
}
int ngp_encoded_mlp_inference(ngp_encoded_mlp_t m, void* stream, const void* params, const void* x, uint32_t x_stride, uint32_t count, void* out) {
    NGP_API_BEGIN require_device(); NGP_CHECK(m && params && x && out, "null argument");
    m->net.inference(S(stream), H(params), F(x), x_stride, count, H(out));
    NGP_API_END
}
int ngp_encoded_mlp_forward_backward(ngp_encoded_mlp_t m, void* stream, const void* params, const void* x, uint32_t x_stride, uint32_t count,
                                     const void* dL, void* grads) {
    NGP_API_BEGIN require_device(); NGP_CHECK(m && params && x && dL && grads, "null argument");
    m->net.forward_backward(S(stream), H(params), F(x), x_stride, count, H(dL), H(grads));
    NGP_API_END
}

// ---- optimizer ----
int ngp_adam_step(void* stream, const ngp_adam_hyper* h, uint64_t n, void* master, void* params, const void* grads, void* m, void* v, void* steps,
                  void* ema, void* ema_half) {
    NGP_API_BEGIN require_device(); NGP_CHECK(h && master && params && grads && m && v && steps, "null argument");
    AdamHyper a{h->learning_rate, h->beta1, h->beta2, h->epsilon, h->l2_reg, h->loss_scale, h->n_matrix_params,
                h->optimize_matrix_params, h->optimize_non_matrix_params, h->ema_decay, h->ema_step};
    adam_step(S(stream), a, n, F(master), H(params), H(grads), F(m), F(v), reinterpret_cast<uint32_t*>(steps), F(ema), H(ema_half));
    NGP_API_END
}

}  // extern "C"

// ---- Testbed ----
struct ngp_testbed { Testbed tb; explicit ngp_testbed(int mode) : tb(mode) {} };
// every entry point sees the testbed with its last step finished on the host side (Testbed::finish_pending_step), except
// ngp_testbed_set_training_image: the upload between two steps must not make the host wait for the device
static inline Testbed& TB(ngp_testbed_t t) { t->tb.finish_pending_step(); return t->tb; }

extern "C" {

int ngp_testbed_create(int mode, ngp_testbed_t* out) {
    NGP_API_BEGIN NGP_CHECK(out, "null argument");
    *out = new ngp_testbed(mode);
    NGP_API_END
}
void ngp_testbed_destroy(ngp_testbed_t t) { delete t; }
int ngp_testbed_reload_network_from_json(ngp_testbed_t t, const char* json) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && json, "null argument");
    TB(t).reload_network_from_json(json);
    NGP_API_END
}
int ngp_testbed_reload_network_from_json_base(ngp_testbed_t t, const char* json, const char* config_base_path) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && json, "null argument");
    TB(t).reload_network_from_json(json, config_base_path ? config_base_path : "");
    NGP_API_END
}
int ngp_testbed_reload_network_from_file(ngp_testbed_t t, const char* path) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    TB(t).reload_network_from_file(path ? path : "");
    NGP_API_END
}
int ngp_testbed_set_root_dir(ngp_testbed_t t, const char* root_dir) {
    NGP_API_BEGIN NGP_CHECK(t && root_dir, "null argument");
    TB(t).root_dir = root_dir;
    NGP_API_END
}
int ngp_builtin_base_config(int32_t mode, const char** json) {
    NGP_API_BEGIN NGP_CHECK(json && mode >= 0 && mode <= 2, "bad argument");
    *json = ngp::Testbed::builtin_base_config(mode);
    NGP_API_END
}
int ngp_testbed_set_nerf_dataset(ngp_testbed_t t, void* stream, uint32_t n_images, const ngp_training_image* images, int32_t aabb_scale) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && images, "null argument");
    static_assert(sizeof(ngp_training_image) == sizeof(HostImage), "ABI struct mismatch");
    TB(t).set_nerf_dataset(S(stream), n_images, reinterpret_cast<const HostImage*>(images), aabb_scale);
    NGP_API_END
}
int ngp_nerf_dataset_open(const char* json_paths, ngp_nerf_dataset_t* out) {
    NGP_API_BEGIN NGP_CHECK(json_paths && out, "null argument");
    std::vector<std::string> paths; std::string cur;
    for (const char* c = json_paths;; ++c) { if (*c == '\n' || *c == 0) { if (!cur.empty()) paths.push_back(cur); cur.clear(); if (!*c) break; } else cur += *c; }
    *out = new ngp_nerf_dataset(paths);
    NGP_API_END
}
void ngp_nerf_dataset_close(ngp_nerf_dataset_t d) { delete d; }
uint32_t ngp_nerf_dataset_n_images(ngp_nerf_dataset_t d) { return d ? (uint32_t)d->ds.n_images() : 0; }
int ngp_nerf_dataset_image_path(ngp_nerf_dataset_t d, uint32_t i, char* buf, uint64_t buf_size) {
    NGP_API_BEGIN NGP_CHECK(d && buf, "null argument"); NGP_CHECK(i < d->ds.n_images(), "Invalid frame index");
    const std::string& p = d->ds.frames[i].resolved_path;
    NGP_CHECK(p.size() + 1 <= buf_size, "image path buffer too small");
    memcpy(buf, p.c_str(), p.size() + 1);
    NGP_API_END
}
int ngp_nerf_dataset_depth_path(ngp_nerf_dataset_t d, uint32_t i, char* buf, uint64_t buf_size) {
    NGP_API_BEGIN NGP_CHECK(d && buf && buf_size > 0, "null argument"); NGP_CHECK(i < d->ds.n_images(), "Invalid frame index");
    const std::string& p = d->ds.frames[i].depth_path;
    NGP_CHECK(p.size() + 1 <= buf_size, "depth_path: buffer too small");
    memcpy(buf, p.c_str(), p.size() + 1);
    NGP_API_END
}
int ngp_nerf_dataset_set_depth(ngp_nerf_dataset_t d, uint32_t i, const uint16_t* pixels, int32_t w, int32_t h) {
    NGP_API_BEGIN NGP_CHECK(d, "null argument");
    d->ds.set_depth(i, pixels, w, h);
    NGP_API_END
}
int ngp_nerf_dataset_set_image(ngp_nerf_dataset_t d, uint32_t i, const void* pixels, int32_t data_type, int32_t w, int32_t h) {
    NGP_API_BEGIN NGP_CHECK(d, "null argument");
    d->ds.set_image(i, pixels, data_type, w, h);
    NGP_API_END
}
int ngp_nerf_dataset_info(ngp_nerf_dataset_t d, int32_t* aabb_scale, float* scale, float* offset3, int32_t* from_mitsuba) {
    NGP_API_BEGIN NGP_CHECK(d, "null argument");
    if (aabb_scale) *aabb_scale = d->ds.aabb_scale; if (scale) *scale = d->ds.scale;
    if (offset3) memcpy(offset3, d->ds.offset, 12); if (from_mitsuba) *from_mitsuba = d->ds.from_mitsuba;
    NGP_API_END
}
static void fill_view(const DatasetFrame& f, ngp_training_image* o) {
    o->pixels = f.pixels.empty() ? nullptr : f.pixels.data(); o->data_type = f.data_type; o->res[0] = f.res[0]; o->res[1] = f.res[1];
    memcpy(o->focal_length, f.focal_length, 8); memcpy(o->principal_point, f.principal_point, 8);
    o->lens_mode = f.lens_mode; memcpy(o->lens_params, f.lens_params, 16); memcpy(o->xform, f.xform, 48);
}
int ngp_nerf_dataset_view(ngp_nerf_dataset_t d, uint32_t i, ngp_training_image* out) {
    NGP_API_BEGIN NGP_CHECK(d && out, "null argument"); NGP_CHECK(i < d->ds.n_images(), "Invalid frame index");
    fill_view(d->ds.frames[i], out);
    NGP_API_END
}
int ngp_nerf_dataset_view_motion(ngp_nerf_dataset_t d, uint32_t i, float* xform_end12, float* rolling_shutter4) {
    NGP_API_BEGIN NGP_CHECK(d && xform_end12 && rolling_shutter4, "null argument"); NGP_CHECK(i < d->ds.n_images(), "Invalid frame index");
    memcpy(xform_end12, d->ds.frames[i].xform_end, 48); memcpy(rolling_shutter4, d->ds.frames[i].rolling_shutter, 16);
    NGP_API_END
}
int ngp_testbed_load_nerf_dataset(ngp_testbed_t t, void* stream, ngp_nerf_dataset_t d) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && d, "null argument");
    NGP_CHECK(d->ds.complete(), "load_nerf_dataset: every frame needs its decoded pixels first (ngp_nerf_dataset_set_image)");
    std::vector<ngp_training_image> views(d->ds.n_images());
    for (size_t i = 0; i < views.size(); ++i) fill_view(d->ds.frames[i], &views[i]);
    static_assert(sizeof(ngp_training_image) == sizeof(HostImage), "ABI struct drift");
    TB(t).set_nerf_dataset(S(stream), (uint32_t)views.size(), reinterpret_cast<const HostImage*>(views.data()), d->ds.aabb_scale);
    TB(t).set_dataset_scale_offset(d->ds.scale, d->ds.offset);
    TB(t).apply_dataset_view_hints(d->ds.has_render_aabb, d->ds.render_aabb, d->ds.up);
    for (size_t i = 0; i < d->ds.frames.size(); ++i) {   // end poses and rolling shutter (already in the NGP convention)
        const ngp::DatasetFrame& f = d->ds.frames[i];
        const bool moving = memcmp(f.xform, f.xform_end, sizeof(f.xform)) != 0;
        const bool shutter = f.rolling_shutter[0] != 0.f || f.rolling_shutter[1] != 0.f || f.rolling_shutter[2] != 0.f || f.rolling_shutter[3] != 0.f;
        if (moving || shutter) TB(t).set_camera_extrinsics_rolling_shutter(S(stream), (int)i, f.xform, f.xform_end, f.rolling_shutter, false);
        // depth maps in scene units: integer_depth_scale * scene scale (nerf_loader.cu:711); a data set with a depth scale gives EVERY view a map
        // (zeros where a frame names no depth file)
        const float depth_scale = f.depth_scale * d->ds.scale;
        if (depth_scale >= 0.f) TB(t).set_training_depth(S(stream), (int)i, f.depth_pixels.empty() ? nullptr : f.depth_pixels.data(), 0, depth_scale);
    }
    NGP_API_END
}
int ngp_testbed_set_training_image(ngp_testbed_t t, void* stream, uint32_t frame_idx, const void* pixels_host) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    t->tb.set_training_image(S(stream), frame_idx, pixels_host);
    NGP_API_END
}
int ngp_testbed_set_camera_intrinsics(ngp_testbed_t t, void* stream, int32_t frame_idx, float fx, float fy, float cx, float cy,
                                      float k1, float k2, float p1, float p2) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    TB(t).set_camera_intrinsics(S(stream), frame_idx, fx, fy, cx, cy, k1, k2, p1, p2);
    NGP_API_END
}
int ngp_testbed_set_camera_extrinsics_rolling_shutter(ngp_testbed_t t, void* stream, int32_t frame_idx, const float* c2w_start, const float* c2w_end,
                                                      const float* rolling_shutter4, int32_t convert_to_ngp) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && c2w_start && c2w_end && rolling_shutter4, "null argument");
    TB(t).set_camera_extrinsics_rolling_shutter(S(stream), frame_idx, c2w_start, c2w_end, rolling_shutter4, convert_to_ngp != 0);
    NGP_API_END
}
int ngp_testbed_set_camera_extrinsics(ngp_testbed_t t, void* stream, int32_t frame_idx, const float* c2w, int32_t convert_to_ngp) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && c2w, "null argument");
    TB(t).set_camera_extrinsics(S(stream), frame_idx, c2w, convert_to_ngp != 0);
    NGP_API_END
}
int ngp_testbed_get_camera_extrinsics(ngp_testbed_t t, int32_t frame_idx, int32_t convert_to_ngp, float* out12) {
    NGP_API_BEGIN NGP_CHECK(t && out12, "null argument");
    TB(t).get_camera_extrinsics(frame_idx, convert_to_ngp != 0, out12);
    NGP_API_END
}
int ngp_testbed_set_training_depth(ngp_testbed_t t, void* stream, int32_t frame_idx, const void* host_depth, int32_t depth_type, float depth_scale) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    TB(t).set_training_depth(S(stream), frame_idx, host_depth, depth_type, depth_scale);
    NGP_API_END
}
int ngp_testbed_reset(ngp_testbed_t t, int32_t reset_density_grid) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    TB(t).reset(reset_density_grid != 0);
    NGP_API_END
}
int ngp_testbed_set_image(ngp_testbed_t t, void* stream, const float* rgba, int32_t w, int32_t h) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && rgba, "null argument");
    TB(t).set_image(S(stream), rgba, w, h);
    NGP_API_END
}
int ngp_testbed_set_image_rgba8(ngp_testbed_t t, void* stream, const uint8_t* rgba8, int32_t w, int32_t h) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && rgba8, "null argument");
    TB(t).set_image_rgba8(S(stream), rgba8, w, h);
    NGP_API_END
}
int ngp_mode_from_scene(const char* path, int32_t* mode) {
    NGP_API_BEGIN NGP_CHECK(path && mode, "null argument");
    struct stat st;
    std::string p = path;
    while (p.size() > 1 && (p.back() == '/')) p.pop_back();
    if (stat(p.c_str(), &st) != 0) { *mode = 4; }                     // ETestbedMode::None: the path does not exist
    else {
        const size_t slash = p.find_last_of('/'), dot = p.find_last_of('.');
        std::string ext = (dot != std::string::npos && (slash == std::string::npos || dot > slash)) ? p.substr(dot + 1) : "";
        for (auto& c : ext) c = (char)tolower((unsigned char)c);
        if (S_ISDIR(st.st_mode) || ext == "json") *mode = 0;
        else if (ext == "obj" || ext == "stl") *mode = 1;
        else if (ext == "nvdb") *mode = 3;
        else *mode = 2;
    }
    NGP_API_END
}
int ngp_snapshot_mode(const char* path, int32_t* mode) {
    NGP_API_BEGIN NGP_CHECK(path && mode, "null argument");
    *mode = ngp::Testbed::snapshot_mode(path);
    NGP_API_END
}
int ngp_testbed_override_sdf_training_data(ngp_testbed_t t, void* stream, const float* points, const float* distances, uint32_t n,
                                           const float* raw_aabb_min, const float* raw_aabb_max, float mesh_scale) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && raw_aabb_min && raw_aabb_max, "null argument");
    TB(t).override_sdf_training_data(S(stream), points, distances, n, raw_aabb_min, raw_aabb_max, mesh_scale);
    NGP_API_END
}
int ngp_testbed_train(ngp_testbed_t t, void* stream, uint32_t batch) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    TB(t).train(S(stream), batch);
    NGP_API_END
}
int ngp_testbed_train_begin(ngp_testbed_t t, void* stream, uint32_t batch) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    TB(t).train_begin(S(stream), batch);
    NGP_API_END
}
int ngp_testbed_train_end(ngp_testbed_t t, void* stream, uint32_t measured, uint32_t measured_before) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t, "null argument");
    TB(t).train_end(S(stream), measured, measured_before);
    NGP_API_END
}
int ngp_testbed_pending_counters(ngp_testbed_t t, uint32_t* measured, uint32_t* measured_before) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && measured && measured_before, "null argument");
    TB(t).pending_counters(measured, measured_before);
    NGP_API_END
}
int ngp_comm_unique_id(void* out128, const char* nccl_lib_path) {
    NGP_API_BEGIN NGP_CHECK(out128, "null argument");
    Comm::unique_id(out128, nccl_lib_path);
    NGP_API_END
}
int ngp_testbed_comm_init(ngp_testbed_t t, const void* id128, uint32_t rank, uint32_t world, const char* nccl_lib_path) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && id128, "null argument");
    TB(t).comm_init(id128, rank, world, nccl_lib_path);
    NGP_API_END
}
int ngp_testbed_set_shard(ngp_testbed_t t, uint32_t rank, uint32_t world) {
    NGP_API_BEGIN NGP_CHECK(t && world >= 1 && rank < world, "bad shard");
    TB(t).set_shard(rank, world);
    NGP_API_END
}
int ngp_testbed_stats(ngp_testbed_t t, ngp_train_stats* out) {
    NGP_API_BEGIN NGP_CHECK(t && out, "null argument");
    static_assert(sizeof(ngp_train_stats) == sizeof(TrainStats), "ABI struct mismatch");
    TrainStats st = TB(t).stats(); memcpy(out, &st, sizeof(st));
    NGP_API_END
}
uint64_t ngp_testbed_n_params(ngp_testbed_t t) { return t ? TB(t).n_params() : 0; }
uint64_t ngp_testbed_n_encoding_params(ngp_testbed_t t) { Trainer* tr = t ? TB(t).trainer() : nullptr; return tr ? tr->n_params - tr->n_matrix_params : 0; }
int ngp_json_to_msgpack(const char* json, void* out, uint64_t capacity, uint64_t* out_size) {
    NGP_API_BEGIN NGP_CHECK(json && out_size, "null argument");
    std::string bytes; msgpack_encode(Json::parse(json), bytes);
    *out_size = bytes.size();
    NGP_CHECK(out && bytes.size() <= capacity, "ngp_json_to_msgpack: output buffer too small");
    memcpy(out, bytes.data(), bytes.size());
    NGP_API_END
}
int ngp_msgpack_roundtrip(const void* in, uint64_t in_size, void* out, uint64_t capacity, uint64_t* out_size) {
    NGP_API_BEGIN NGP_CHECK(in && out_size, "null argument");
    std::string bytes; msgpack_encode(msgpack_decode(in, in_size), bytes);
    *out_size = bytes.size();
    NGP_CHECK(out && bytes.size() <= capacity, "ngp_msgpack_roundtrip: output buffer too small");
    memcpy(out, bytes.data(), bytes.size());
    NGP_API_END
}
int ngp_testbed_save_snapshot(ngp_testbed_t t, void* stream, const char* path, int32_t include_optimizer_state, int32_t compress) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && path, "null argument");
    TB(t).save_snapshot(S(stream), path, include_optimizer_state != 0, compress != 0);
    NGP_API_END
}
int ngp_testbed_load_snapshot(ngp_testbed_t t, void* stream, const char* path) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && path, "null argument");
    TB(t).load_snapshot(S(stream), path);
    NGP_API_END
}
int ngp_testbed_sync_inference_params(ngp_testbed_t t, void* stream) {
    NGP_API_BEGIN NGP_CHECK(t, "null argument"); require_device();
    TB(t).sync_inference_params(S(stream));
    NGP_API_END
}
int ngp_testbed_param_buffer(ngp_testbed_t t, int which, void** ptr, uint64_t* n) {
    NGP_API_BEGIN NGP_CHECK(t && ptr && n, "null argument");
    Trainer* tr = TB(t).trainer(); NGP_CHECK(tr, "no network loaded");
    switch (which) { case 0: *ptr = tr->master; break; case 1: *ptr = tr->params; break; case 2: *ptr = tr->inference_params; break;
        case 3: *ptr = tr->grads; break; case 4: *ptr = tr->m; break; case 5: *ptr = tr->v; break; case 6: *ptr = tr->steps; break;
        case 7: *ptr = tr->ema; break; default: throw Error("param_buffer: which must be 0..7"); }
    *n = tr->n_params;
    NGP_API_END
}
int ngp_testbed_density_grid(ngp_testbed_t t, int which, void** ptr, uint64_t* n_bytes) {
    NGP_API_BEGIN NGP_CHECK(t && ptr && n_bytes, "null argument");
    if (which == 0) { *ptr = (void*)TB(t).density_grid(); *n_bytes = (uint64_t)NERF_GRID_N_CELLS * (TB(t).max_cascade() + 1) * 4; }
    else if (which == 1) { *ptr = (void*)TB(t).density_grid_bitfield(); *n_bytes = (uint64_t)NERF_GRID_N_CELLS * NERF_CASCADES / 8; }
    else throw Error("density_grid: which must be 0 or 1");
    NGP_API_END
}
int ngp_testbed_render(ngp_testbed_t t, void* stream, const float* cam, const float* focal, const float* center, int32_t w, int32_t h, int32_t spp,
                       int32_t y0, int32_t y1, float* out, int32_t out_is_device, float* depth) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && cam && focal && center && out, "null argument");
    TB(t).render(S(stream), cam, focal, center, w, h, spp, y0, y1, out, out_is_device != 0, depth);
    NGP_API_END
}
int ngp_testbed_render_frame(ngp_testbed_t t, void* stream, int32_t w, int32_t h, int32_t spp, int32_t linear, int32_t y0, int32_t y1, void* out, int32_t out_is_device) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && out, "null argument");
    TB(t).render_frame(S(stream), w, h, spp, linear == 0, y0, y1, reinterpret_cast<float*>(out), out_is_device != 0);
    NGP_API_END
}
int ngp_testbed_reset_camera(ngp_testbed_t t) { NGP_API_BEGIN NGP_CHECK(t, "null argument"); TB(t).reset_camera(); NGP_API_END }
int ngp_testbed_set_nerf_camera_matrix(ngp_testbed_t t, const float* m) { NGP_API_BEGIN NGP_CHECK(t && m, "null argument"); TB(t).set_nerf_camera_matrix(m); NGP_API_END }
int ngp_testbed_set_camera_to_training_view(ngp_testbed_t t, int32_t v) { NGP_API_BEGIN NGP_CHECK(t, "null argument"); TB(t).set_camera_to_training_view(v); NGP_API_END }
namespace {
float* vector_field(ngp::Testbed& tb, const std::string& n, uint32_t count) {
    struct E { const char* name; float* p; uint32_t n; };
    const E table[] = {{"camera_matrix", tb.camera, 12}, {"screen_center", tb.screen_center, 2}, {"relative_focal_length", tb.relative_focal_length, 2},
                       {"up_dir", tb.up_dir, 3}, {"background_color", tb.background_color, 4}, {"aabb", tb.aabb.min, 6}, {"render_aabb", tb.render_aabb.min, 6}};
    for (const E& e : table) if (n == e.name) { NGP_CHECK(count == e.n, "vector option '" + n + "' has " + std::to_string(e.n) + " components"); return e.p; }
    throw ngp::Error("unknown vector option '" + n + "'");
}
}
int ngp_testbed_set_vector(ngp_testbed_t t, const char* name, const float* v, uint32_t count) {
    NGP_API_BEGIN NGP_CHECK(t && name && v, "null argument");
    memcpy(vector_field(t->tb, name, count), v, count * sizeof(float));
    NGP_API_END
}
int ngp_testbed_get_vector(ngp_testbed_t t, const char* name, float* v, uint32_t count) {
    NGP_API_BEGIN NGP_CHECK(t && name && v, "null argument");
    if (std::string(name) == "dataset_offset") { NGP_CHECK(count == 3, "vector option 'dataset_offset' has 3 components"); memcpy(v, TB(t).dataset_offset(), 12); return 0; }   // read-only
    memcpy(v, vector_field(t->tb, name, count), count * sizeof(float));
    NGP_API_END
}
int ngp_testbed_set_render_lens(ngp_testbed_t t, int32_t lens_mode, const float* p) {
    NGP_API_BEGIN NGP_CHECK(t, "null argument");
    NGP_CHECK(lens_mode == LENS_PERSPECTIVE || (lens_mode == LENS_OPENCV && p), "render lens must be Perspective (0) or OpenCV (1) with 4 parameters");
    TB(t).render_lens_mode = lens_mode;
    for (int k = 0; k < 4; ++k) TB(t).render_lens_params[k] = (lens_mode == LENS_OPENCV) ? p[k] : 0.f;
    NGP_API_END
}
uint32_t ngp_testbed_n_training_views(ngp_testbed_t t) { return t ? (uint32_t)TB(t).training_views().size() : 0; }
int ngp_testbed_training_view(ngp_testbed_t t, uint32_t i, ngp_training_image* o) {
    NGP_API_BEGIN NGP_CHECK(t && o, "null argument"); NGP_CHECK(i < TB(t).training_views().size(), "Invalid frame index");
    const ImageMeta& m = TB(t).training_views()[i];
    o->pixels = m.pixels; o->data_type = m.data_type; o->res[0] = m.res[0]; o->res[1] = m.res[1];
    memcpy(o->focal_length, m.focal_length, 8); memcpy(o->principal_point, m.principal_point, 8);
    o->lens_mode = m.lens_mode; memcpy(o->lens_params, m.lens_params, 16); memcpy(o->xform, m.xform, 48);
    NGP_API_END
}
int ngp_testbed_training_view_motion(ngp_testbed_t t, uint32_t i, float* xform_start12, float* xform_end12, float* rolling_shutter4) {
    NGP_API_BEGIN NGP_CHECK(t, "null argument"); NGP_CHECK(i < TB(t).training_views().size() && TB(t).given_xform(i), "Invalid frame index");
    if (xform_start12) memcpy(xform_start12, TB(t).given_xform(i), 48);
    if (xform_end12) memcpy(xform_end12, TB(t).given_xform(i) + 12, 48);
    if (rolling_shutter4) memcpy(rolling_shutter4, TB(t).training_views()[i].rolling_shutter, 16);
    NGP_API_END
}
int ngp_testbed_compute_image_mse(ngp_testbed_t t, void* stream, int32_t quantize, float* out) {
    NGP_API_BEGIN require_device(); NGP_CHECK(t && out, "null argument");
    *out = TB(t).compute_image_mse(S(stream), quantize != 0);
    NGP_API_END
}

int ngp_testbed_debug_buffer(ngp_testbed_t t, const char* name, void** ptr, uint64_t* n_bytes) {
    NGP_API_BEGIN NGP_CHECK(t && name && ptr && n_bytes, "null argument");
    TB(t).debug_buffer(name, ptr, n_bytes);
    NGP_API_END
}

#define NGP_OPTIONS(X)                                                                                                           \
    X(random_bg_color, bool) X(linear_colors, bool) X(snap_to_pixel_centers, bool) X(train_snap_to_pixel_centers, bool) X(shall_train_encoding, bool) X(shall_train_network, bool) \
    X(near_distance, float) X(density_grid_decay, float) X(render_min_transmittance, float) X(cone_angle_constant, float)         \
    X(loss_type, int) X(rgb_activation, int) X(density_activation, int) X(seed, uint64_t)                                        \
    X(image_snap_to_pixel_centers, bool) X(image_linear_colors, bool) X(image_random_mode, int) X(sampler_lanes_per_ray, uint32_t) X(prefetch_rays, bool) X(prefetch_hits, uint32_t) X(dp_sharded_optimizer, bool) X(device_control_law, bool) \
    X(exposure, float) X(zoom, float) X(scale, float) X(ground_truth_alpha, float) X(render_near_distance, float) X(fov_axis, int) X(training_view, int)                      \
    X(color_space, int) X(tonemap_curve, int) X(render_mode, int) X(render_ground_truth, bool) X(render_with_lens_distortion, bool) X(render_row_stride, int) X(walk_ahead, bool) X(walk_ahead_hits, uint32_t) X(walk_ahead_misses, uint32_t) \
    X(optimize_exposure, bool) X(exposure_l2_reg, float) X(n_steps_between_cam_updates, uint32_t) X(max_level_rand_training, bool) X(depth_supervision_lambda, float) X(depth_loss_type, int) X(deferred_step_finish, bool) X(optimize_extrinsics, bool) X(optimize_distortion, bool) X(extrinsic_l2_reg, float) X(extrinsic_learning_rate, float) \
    X(sample_focal_plane_proportional_to_error, bool) X(sample_image_proportional_to_error, bool) X(n_steps_between_error_map_updates, uint32_t)

int ngp_testbed_set_option(ngp_testbed_t t, const char* name, double value) {
    NGP_API_BEGIN NGP_CHECK(t && name, "null argument");
    const std::string n = name;
#define X(field, type) if (n == #field) { TB(t).field = (type)value; return 0; }
    NGP_OPTIONS(X)
#undef X
    if (n == "background_r") { TB(t).background_color[0] = (float)value; return 0; }
    if (n == "background_g") { TB(t).background_color[1] = (float)value; return 0; }
    if (n == "background_b") { TB(t).background_color[2] = (float)value; return 0; }
    if (n == "background_a") { TB(t).background_color[3] = (float)value; return 0; }
    if (n == "fov") { TB(t).set_fov((float)value); return 0; }
    if (n == "image_stratified") { TB(t).image_random_mode = value != 0 ? IMG_RANDOM_STRATIFIED : IMG_RANDOM_RANDOM; return 0; }   // round-1 name
    throw Error("unknown option '" + n + "'");
    NGP_API_END
}
int ngp_testbed_get_option(ngp_testbed_t t, const char* name, double* value) {
    NGP_API_BEGIN NGP_CHECK(t && name && value, "null argument");
    const std::string n = name;
    if (n == "training_step") { *value = (double)t->tb.training_step(); return 0; }   // read-only; does not wait for a pending step's counters (run.py:183-202 polls it every frame)
#define X(field, type) if (n == #field) { *value = (double)TB(t).field; return 0; }
    NGP_OPTIONS(X)
#undef X
    if (n == "background_r") { *value = TB(t).background_color[0]; return 0; }
    if (n == "background_g") { *value = TB(t).background_color[1]; return 0; }
    if (n == "background_b") { *value = TB(t).background_color[2]; return 0; }
    if (n == "background_a") { *value = TB(t).background_color[3]; return 0; }
    if (n == "fov") { *value = TB(t).fov(); return 0; }
    if (n == "image_stratified") { *value = TB(t).image_random_mode == IMG_RANDOM_STRATIFIED; return 0; }
    if (n == "error_map_width") { *value = TB(t).error_map_res()[0]; return 0; }            // read-only: the lattice of the current error-map window
    if (n == "error_map_height") { *value = TB(t).error_map_res()[1]; return 0; }
    if (n == "error_cdf_width") { *value = TB(t).error_cdf_res()[0]; return 0; }            // ... and of the CDFs in use (0: none yet)
    if (n == "error_cdf_height") { *value = TB(t).error_cdf_res()[1]; return 0; }
    if (n == "dataset_scale") { *value = TB(t).dataset_scale(); return 0; }              // read-only: nerf.training.dataset.scale / aabb_scale
    if (n == "dataset_aabb_scale") { *value = TB(t).dataset_aabb_scale(); return 0; }
    if (n == "measured_batch_size_total") { *value = (double)TB(t).measured_total(); return 0; }                          // read-only: sums over all steps so far
    if (n == "measured_batch_size_before_compaction_total") { *value = (double)TB(t).measured_before_total(); return 0; }
    if (n == "l2_window_bytes") { *value = TB(t).trainer() ? (double)TB(t).trainer()->l2_window_bytes : 0.0; return 0; }   // read-only
    throw Error("unknown option '" + n + "'");
    NGP_API_END
}

}  // extern "C"
