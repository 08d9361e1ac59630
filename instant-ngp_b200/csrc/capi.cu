// extern "C" boundary (include/ngp_b200.h): operator-level entry points.  Exceptions never cross the ABI — they are
// caught, stored per thread and reported as a non-zero status + ngp_last_error().
#include "../../include/ngp_b200.h"
#include "network.h"

namespace ngp { std::atomic<uint64_t> g_launch_count{0}; }

static thread_local std::string g_last_error;

#define NGP_API_BEGIN try {
#define NGP_API_END                                            \
    return 0;                                                  \
    } catch (const std::exception& e) { g_last_error = e.what(); return 1; } \
      catch (...) { g_last_error = "unknown error"; return 1; }

using namespace ngp;

struct ngp_encoding { GridLayout g; };
struct ngp_nerf_network { NerfNetwork net; explicit ngp_nerf_network(const Json& j) : net(j) {} };
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

extern "C" {

const char* ngp_last_error(void) { return g_last_error.c_str(); }
int ngp_abi_version(void) { return NGP_B200_ABI_VERSION; }
uint64_t ngp_kernel_launch_count(void) { return g_launch_count.load(); }

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
    NGP_API_BEGIN require_device(); NGP_CHECK(e && params && x && y, "null argument");
    grid_forward(S(stream), e->g, H(params), F(x), x_stride, n, H(y), y_layout);
    NGP_API_END
}
int ngp_encoding_backward(ngp_encoding_t e, void* stream, const void* x, uint32_t x_stride, uint32_t n, const void* dy, int dy_layout, void* grad) {
    NGP_API_BEGIN require_device(); NGP_CHECK(e && x && dy && grad, "null argument");
    grid_backward(S(stream), e->g, F(x), x_stride, n, H(dy), dy_layout, H(grad));
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
