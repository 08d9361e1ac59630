// ORACLE — TEST INFRASTRUCTURE ONLY (see ngp_ref.h). Core operators: pcg32, hash grid, SH4, MLP,
// NerfNetwork wiring, Adam/EMA/ExponentialDecay.  Parity unpinned for the tcnn-resident pieces.
#include "ngp_ref.h"
#include <cmath>
#include <cstring>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

// ---------------------------------------------------------------------------------------------
// fp16 storage emulation
// ---------------------------------------------------------------------------------------------
static inline float h2f(uint16_t h) { _Float16 v; memcpy(&v, &h, 2); return (float)v; }
static inline uint16_t f2h(float f) { _Float16 v = (_Float16)f; uint16_t h; memcpy(&h, &v, 2); return h; }
static inline float rh(float f) { return h2f(f2h(f)); }   // round through fp16

extern "C" void ref_float_to_half(const float* in, uint16_t* out, size_t n) { for (size_t i = 0; i < n; ++i) out[i] = f2h(in[i]); }
extern "C" void ref_half_to_float(const uint16_t* in, float* out, size_t n) { for (size_t i = 0; i < n; ++i) out[i] = h2f(in[i]); }

// ---------------------------------------------------------------------------------------------
// pcg32 — tcnn::default_rng_t (used at testbed.h:609 seed 1337, testbed_nerf.cu:1421 advance(i*16), ...)
// Published algorithm: M. O'Neill, "PCG: A Family of Simple Fast Space-Efficient ..." (pcg32 XSH-RR 64/32).
// ---------------------------------------------------------------------------------------------
static const uint64_t PCG32_MULT = 0x5851f42d4c957f2dULL;

extern "C" uint32_t ref_pcg32_next_uint(RefPcg32* r) {
    uint64_t old = r->state;
    r->state = old * PCG32_MULT + r->inc;
    uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
}
extern "C" void ref_pcg32_seed(RefPcg32* r, uint64_t initstate, uint64_t initseq) {
    r->state = 0u;
    r->inc = (initseq << 1u) | 1u;
    ref_pcg32_next_uint(r);
    r->state += initstate;
    ref_pcg32_next_uint(r);
}
extern "C" float ref_pcg32_next_float(RefPcg32* r) {
    uint32_t u = (ref_pcg32_next_uint(r) >> 9) | 0x3f800000u;
    float f; memcpy(&f, &u, 4);
    return f - 1.0f;
}
extern "C" void ref_pcg32_advance(RefPcg32* r, int64_t delta_) {
    uint64_t cur_mult = PCG32_MULT, cur_plus = r->inc, acc_mult = 1u, acc_plus = 0u;
    uint64_t delta = (uint64_t)delta_;
    while (delta > 0) {
        if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta /= 2;
    }
    r->state = acc_mult * r->state + acc_plus;
}
extern "C" void ref_pcg32_fill_float(uint64_t seed, uint64_t skip, float* out, size_t n) {
    RefPcg32 r; ref_pcg32_seed(&r, seed, 1); ref_pcg32_advance(&r, (int64_t)skip);
    for (size_t i = 0; i < n; ++i) out[i] = ref_pcg32_next_float(&r);
}
extern "C" void ref_pcg32_fill_uint(uint64_t seed, uint64_t skip, uint32_t* out, size_t n) {
    RefPcg32 r; ref_pcg32_seed(&r, seed, 1); ref_pcg32_advance(&r, (int64_t)skip);
    for (size_t i = 0; i < n; ++i) out[i] = ref_pcg32_next_uint(&r);
}

// ---------------------------------------------------------------------------------------------
// Hash grid (tcnn GridEncoding; call sites nerf_network.h:93-95,143-148,201-207,324-333,345-350;
// config derivation testbed.cu:3961-4005).  Level rule recalled from tcnn:
//   scale_l = 2^(l*log2 b) * N_min - 1 ; res_l = ceil(scale_l) + 1
//   size_l  = min(round_up(res_l^D, 8), T)   (dense if res^D <= T else hashed)
//   pos = x*scale + 0.5 ; cell = floor(pos) ; w = pos - cell ; 2^D corners, D-linear weights
//   index = sum_d cell_d * res^d   (while stride <= size) ; if size < stride: xor_d cell_d*prime_d ; % size
// ---------------------------------------------------------------------------------------------
struct Level { uint32_t offset, size, res; float scale; };

static void grid_levels(const RefGridCfg* c, std::vector<Level>& lv) {
    lv.resize(c->n_levels);
    float log2b = std::log2(c->per_level_scale);
    uint32_t offset = 0;
    for (uint32_t l = 0; l < c->n_levels; ++l) {
        float scale = exp2f((float)l * log2b) * (float)c->base_resolution - 1.0f;
        uint32_t res = (uint32_t)ceilf(scale) + 1u;
        uint32_t max_params = 0xFFFFFFFFu / 2;
        double dense = std::pow((double)res, (double)c->n_dims);
        uint32_t n = dense > (double)max_params ? max_params : (uint32_t)(dense + 0.5);
        n = (n + REF_GRID_PARAM_ALIGN - 1) / REF_GRID_PARAM_ALIGN * REF_GRID_PARAM_ALIGN;
        n = std::min(n, 1u << c->log2_hashmap_size);
        lv[l] = {offset, n, res, scale};
        offset += n;
    }
}

extern "C" uint64_t ref_grid_layout(const RefGridCfg* c, uint32_t* offsets, float* scales, uint32_t* resolutions) {
    std::vector<Level> lv; grid_levels(c, lv);
    uint32_t end = 0;
    for (uint32_t l = 0; l < c->n_levels; ++l) {
        if (offsets) offsets[l] = lv[l].offset;
        if (scales) scales[l] = lv[l].scale;
        if (resolutions) resolutions[l] = lv[l].res;
        end = lv[l].offset + lv[l].size;
    }
    if (offsets) offsets[c->n_levels] = end;
    return (uint64_t)end * c->n_features;
}

static inline uint32_t grid_index(uint32_t D, uint32_t size, uint32_t res, const uint32_t* cell) {
    uint32_t stride = 1, index = 0;
    for (uint32_t d = 0; d < D && stride <= size; ++d) { index += cell[d] * stride; stride *= res; }
    if (size < stride) {
        static const uint32_t primes[3] = {1u, REF_GRID_PRIME_1, REF_GRID_PRIME_2};
        index = 0;
        for (uint32_t d = 0; d < D; ++d) index ^= cell[d] * primes[d];
    }
    return index % size;
}

// corners of one (point, level): idx[2^D], w[2^D]; also returns frac/cell for input gradients
static inline void corners(uint32_t D, const Level& L, const float* x, uint32_t* idx, float* w, float* frac_out = nullptr,
                           uint32_t* cell_out = nullptr) {
    float frac[3]; uint32_t cell[3];
    for (uint32_t d = 0; d < D; ++d) {
        float p = fmaf(L.scale, x[d], 0.5f);
        float f = floorf(p);
        cell[d] = (uint32_t)(int)f;
        frac[d] = p - f;
        if (frac_out) frac_out[d] = frac[d];
        if (cell_out) cell_out[d] = cell[d];
    }
    for (uint32_t k = 0; k < (1u << D); ++k) {
        float weight = 1.0f; uint32_t cl[3];
        for (uint32_t d = 0; d < D; ++d) {
            if ((k & (1u << d)) == 0) { weight *= 1.0f - frac[d]; cl[d] = cell[d]; }
            else { weight *= frac[d]; cl[d] = cell[d] + 1; }
        }
        idx[k] = grid_index(D, L.size, L.res, cl);
        w[k] = weight;
    }
}

extern "C" void ref_grid_corners(const RefGridCfg* c, const float* x, uint32_t xs, uint32_t n, uint32_t level,
                                 uint32_t* idx_out, float* w_out) {
    std::vector<Level> lv; grid_levels(c, lv);
    const uint32_t K = 1u << c->n_dims;
    for (uint32_t i = 0; i < n; ++i) corners(c->n_dims, lv[level], x + (size_t)i * xs, idx_out + (size_t)i * K, w_out + (size_t)i * K);
}

// GridEncoding::set_max_level_gpu (tiny-cuda-nn grid.h kernel_grid / kernel_grid_backward, RECALLED - the source is not in the mount): with a
// per-row cut-off m the forward pass writes zeros for the levels l >= m * n_levels + 1e-3 and the backward pass skips the levels
// l > m * n_levels + 1e-3 (the two comparisons differ in the library).  The pointer applies to the ref_grid_forward / ref_grid_backward calls
// that follow (rows are indexed from the start of x); NULL = every level.
static thread_local const float* g_grid_max_level = nullptr;   // thread-local: the per-sample network functions below set it inside their OpenMP loops
extern "C" void ref_grid_set_max_level(const float* max_level) { g_grid_max_level = max_level; }

extern "C" void ref_grid_forward(const RefGridCfg* c, const uint16_t* table, const float* x, uint32_t xs, uint32_t n,
                                 uint16_t* y, uint32_t ys, uint32_t y_width) {
    std::vector<Level> lv; grid_levels(c, lv);
    const uint32_t D = c->n_dims, F = c->n_features, K = 1u << D;
    const float* max_level = g_grid_max_level;   // the caller's (the loop's threads have their own thread-local)
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        uint16_t* yo = y + (size_t)i * ys;
        for (uint32_t l = 0; l < c->n_levels; ++l) {
            if (max_level && (float)l >= max_level[i] * (float)c->n_levels + 1e-3f) { for (uint32_t f = 0; f < F; ++f) yo[l * F + f] = 0; continue; }
            uint32_t idx[8]; float w[8];
            corners(D, lv[l], x + (size_t)i * xs, idx, w);
            const uint16_t* t = table + (size_t)lv[l].offset * F;
            for (uint32_t f = 0; f < F; ++f) {
                float acc = 0.0f;
                for (uint32_t k = 0; k < K; ++k) acc = fmaf(w[k], h2f(t[(size_t)idx[k] * F + f]), acc);
                yo[l * F + f] = f2h(acc);
            }
        }
        for (uint32_t j = c->n_levels * F; j < y_width; ++j) yo[j] = 0;
    }
}

extern "C" void ref_grid_backward(const RefGridCfg* c, const float* x, uint32_t xs, uint32_t n,
                                  const uint16_t* dy, uint32_t dys, double* grad) {
    std::vector<Level> lv; grid_levels(c, lv);
    const uint32_t D = c->n_dims, F = c->n_features, K = 1u << D;
    for (uint32_t i = 0; i < n; ++i) {
        for (uint32_t l = 0; l < c->n_levels; ++l) {
            if (g_grid_max_level && (float)l > g_grid_max_level[i] * (float)c->n_levels + 1e-3f) continue;
            uint32_t idx[8]; float w[8];
            corners(D, lv[l], x + (size_t)i * xs, idx, w);
            double* g = grad + (size_t)lv[l].offset * F;
            for (uint32_t f = 0; f < F; ++f) {
                float d = h2f(dy[(size_t)i * dys + l * F + f]);
                if (d == 0.0f) continue;
                for (uint32_t k = 0; k < K; ++k) g[(size_t)idx[k] * F + f] += (double)w[k] * (double)d;
            }
        }
    }
}

// dy/dx (tcnn kernel_grid_backward_input; call sites testbed_nerf.cu:2615-2617,4071-4078): linear interpolation,
// d y_f / d x_d = scale * sum over the 2^(D-1) corner pairs of w_others * (val(cell_d+1) - val(cell_d))
extern "C" void ref_grid_backward_input(const RefGridCfg* c, const uint16_t* table, const float* x, uint32_t xs,
                                        uint32_t n, const uint16_t* dy, uint32_t dys, float* dx) {
    std::vector<Level> lv; grid_levels(c, lv);
    const uint32_t D = c->n_dims, F = c->n_features;
    for (uint32_t i = 0; i < n; ++i) {
        float out[3] = {0, 0, 0};
        for (uint32_t l = 0; l < c->n_levels; ++l) {
            uint32_t idx[8], cell[3]; float w[8], frac[3];
            corners(D, lv[l], x + (size_t)i * xs, idx, w, frac, cell);
            const uint16_t* t = table + (size_t)lv[l].offset * F;
            for (uint32_t gd = 0; gd < D; ++gd) {
                float acc = 0.0f;
                for (uint32_t k = 0; k < (1u << D); ++k) {
                    if (k & (1u << gd)) continue;
                    float wo = 1.0f;
                    for (uint32_t d = 0; d < D; ++d) if (d != gd) wo *= (k & (1u << d)) ? frac[d] : 1.0f - frac[d];
                    for (uint32_t f = 0; f < F; ++f) {
                        float v0 = h2f(t[(size_t)idx[k] * F + f]);
                        float v1 = h2f(t[(size_t)idx[k | (1u << gd)] * F + f]);
                        acc += wo * (v1 - v0) * h2f(dy[(size_t)i * dys + l * F + f]);
                    }
                }
                out[gd] += acc * lv[l].scale;
            }
        }
        for (uint32_t d = 0; d < D; ++d) dx[(size_t)i * D + d] = out[d];
    }
}

// init: element i = U(-1e-4,1e-4) from float #i of the stream (counter-based so the GPU fill matches)
static void grid_init_from(const RefGridCfg* c, RefPcg32 base, float* p) {
    uint64_t n = ref_grid_layout(c, nullptr, nullptr, nullptr);
    RefPcg32 r = base;
    for (uint64_t i = 0; i < n; ++i) p[i] = ref_pcg32_next_float(&r) * (2.0f * REF_GRID_INIT_RANGE) - REF_GRID_INIT_RANGE;
}
extern "C" void ref_grid_init(const RefGridCfg* c, uint64_t seed, float* p) {
    RefPcg32 r; ref_pcg32_seed(&r, seed, 1); grid_init_from(c, r, p);
}

// ---------------------------------------------------------------------------------------------
// SphericalHarmonics degree 4 (tcnn; config configs/nerf/base.json:37-49; nerf_network.h:97,155-162).
// Real SH basis of the unit-ish vector 2*d-1, 16 coefficients, fp16 out.
// ---------------------------------------------------------------------------------------------
static inline void sh4_one(const float* d, float* o) {
    float x = d[0] * 2.f - 1.f, y = d[1] * 2.f - 1.f, z = d[2] * 2.f - 1.f;
    float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = 0.28209479177387814f;
    o[1] = -0.48860251190291987f * y;
    o[2] = 0.48860251190291987f * z;
    o[3] = -0.48860251190291987f * x;
    o[4] = 1.0925484305920792f * xy;
    o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz;
    o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
    o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
    o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
    o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}
// d(SH4 outputs)/d(direction input in [0,1]^3): J[k][axis], the analytic partials of the polynomials in sh4_one times 2 (x = 2 d - 1).
// tiny-cuda-nn's SphericalHarmonicsEncoding hands exactly these to its backward pass (dy_dx); restated from the polynomials, pinned by finite
// differences in tests/test_oracle_cpu.py.
static void sh4_jacobian(const float* d, float J[16][3]) {
    const float x = d[0] * 2.f - 1.f, y = d[1] * 2.f - 1.f, z = d[2] * 2.f - 1.f;
    const float C1 = 0.48860251190291987f, C2 = 1.0925484305920792f, C3 = 0.94617469575755997f, C4 = 0.54627421529603959f, C5 = 0.59004358992664352f,
                C6 = 2.8906114426405538f, C7 = 0.45704579946446572f, C8 = 0.3731763325901154f, C9 = 1.4453057213202769f;
    const float P[16][3] = {
        {0.f, 0.f, 0.f},
        {0.f, -C1, 0.f},
        {0.f, 0.f, C1},
        {-C1, 0.f, 0.f},
        {C2 * y, C2 * x, 0.f},
        {0.f, -C2 * z, -C2 * y},
        {0.f, 0.f, 2.f * C3 * z},
        {-C2 * z, 0.f, -C2 * x},
        {2.f * C4 * x, -2.f * C4 * y, 0.f},
        {C5 * y * (-6.f * x), C5 * (-3.f * x * x + 3.f * y * y), 0.f},
        {C6 * y * z, C6 * x * z, C6 * x * y},
        {0.f, C7 * (1.f - 5.f * z * z), C7 * y * (-10.f * z)},
        {0.f, 0.f, C8 * (15.f * z * z - 3.f)},
        {C7 * (1.f - 5.f * z * z), 0.f, C7 * x * (-10.f * z)},
        {C9 * z * 2.f * x, C9 * z * (-2.f * y), C9 * (x * x - y * y)},
        {C5 * (-3.f * x * x + 3.f * y * y), C5 * x * (6.f * y), 0.f}};
    for (int k = 0; k < 16; ++k) for (int a = 0; a < 3; ++a) J[k][a] = 2.f * P[k][a];
}
extern "C" void ref_sh4_jacobian(const float* dir3, float* J48) { float J[16][3]; sh4_jacobian(dir3, J); memcpy(J48, J, sizeof(J)); }

extern "C" void ref_sh4(const float* dirs, uint32_t stride, uint32_t n, uint16_t* out, uint32_t os) {
    for (uint32_t i = 0; i < n; ++i) {
        float o[16]; sh4_one(dirs + (size_t)i * stride, o);
        for (int j = 0; j < 16; ++j) out[(size_t)i * os + j] = f2h(o[j]);
    }
}

// ---------------------------------------------------------------------------------------------
// FullyFusedMLP (tcnn; call sites nerf_network.h:99-111,150-153,164,211-216,236-241,279,315).
// ReLU hidden layers, linear output padded to 16, fp16 weights/activations.  The reference's wmma
// accumulates in fp16; this oracle accumulates in fp32 and rounds at the same storage points —
// the 2e-3 tolerance in BASELINE.json absorbs the difference.
// ---------------------------------------------------------------------------------------------
extern "C" uint32_t ref_mlp_n_params(uint32_t in_w, uint32_t W, uint32_t nh, uint32_t out_pad) {
    return W * in_w + (nh - 1) * W * W + out_pad * W;
}

static void mlp_fwd_one(const uint16_t* wts, uint32_t in_w, uint32_t W, uint32_t nh, uint32_t out_pad,
                        const float* in, float* out, float* acts /* [nh][W] or null */) {
    float cur[128], nxt[128];
    for (uint32_t i = 0; i < in_w; ++i) cur[i] = in[i];
    uint32_t cw = in_w;
    const uint16_t* w = wts;
    for (uint32_t l = 0; l < nh; ++l) {
        for (uint32_t o = 0; o < W; ++o) {
            float acc = 0.f;
            for (uint32_t i = 0; i < cw; ++i) acc = fmaf(h2f(w[o * cw + i]), cur[i], acc);
            nxt[o] = rh(acc > 0.f ? acc : 0.f);
        }
        if (acts) for (uint32_t o = 0; o < W; ++o) acts[l * W + o] = nxt[o];
        for (uint32_t o = 0; o < W; ++o) cur[o] = nxt[o];
        w += (size_t)W * cw; cw = W;
    }
    for (uint32_t o = 0; o < out_pad; ++o) {
        float acc = 0.f;
        for (uint32_t i = 0; i < cw; ++i) acc = fmaf(h2f(w[o * cw + i]), cur[i], acc);
        out[o] = rh(acc);
    }
}

extern "C" void ref_mlp_forward(const uint16_t* wts, uint32_t in_w, uint32_t W, uint32_t nh, uint32_t out_pad,
                                const uint16_t* in, uint32_t is, uint32_t n, uint16_t* out, uint32_t os, uint16_t* acts) {
    #pragma omp parallel for schedule(static)
    for (int64_t s = 0; s < (int64_t)n; ++s) {
        float x[128], y[16], a[8 * 128];
        for (uint32_t i = 0; i < in_w; ++i) x[i] = h2f(in[(size_t)s * is + i]);
        mlp_fwd_one(wts, in_w, W, nh, out_pad, x, y, acts ? a : nullptr);
        for (uint32_t o = 0; o < out_pad; ++o) out[(size_t)s * os + o] = f2h(y[o]);
        if (acts) for (uint32_t l = 0; l < nh; ++l) for (uint32_t o = 0; o < W; ++o)
            acts[((size_t)l * n + s) * W + o] = f2h(a[l * W + o]);
    }
}

// one-sample backward. dout[out_pad] (already fp16-rounded floats). din (optional) [in_w] rounded to fp16.
static void mlp_bwd_one(const uint16_t* wts, uint32_t in_w, uint32_t W, uint32_t nh, uint32_t out_pad,
                        const float* in, const float* acts, const float* dout, float* din, double* wgrad) {
    // layer offsets
    std::vector<size_t> off(nh + 1); std::vector<uint32_t> cin(nh + 1);
    size_t o = 0; uint32_t cw = in_w;
    for (uint32_t l = 0; l < nh; ++l) { off[l] = o; cin[l] = cw; o += (size_t)W * cw; cw = W; }
    off[nh] = o; cin[nh] = W;
    float g[128], gp[128];
    for (uint32_t i = 0; i < out_pad; ++i) g[i] = dout[i];
    uint32_t gw = out_pad;
    for (int l = (int)nh; l >= 0; --l) {
        const float* x = (l == 0) ? in : acts + (size_t)(l - 1) * W;
        uint32_t xw = cin[l];
        const uint16_t* w = wts + off[l];
        if (wgrad) for (uint32_t r = 0; r < gw; ++r) { if (g[r] == 0.f) continue;
            for (uint32_t i = 0; i < xw; ++i) wgrad[off[l] + (size_t)r * xw + i] += (double)g[r] * (double)x[i]; }
        if (l == 0 && !din) break;
        for (uint32_t i = 0; i < xw; ++i) {
            float acc = 0.f;
            for (uint32_t r = 0; r < gw; ++r) acc = fmaf(h2f(w[(size_t)r * xw + i]), g[r], acc);
            if (l > 0) acc = (x[i] > 0.f) ? acc : 0.f;   // ReLU mask from the saved activation
            gp[i] = rh(acc);
        }
        for (uint32_t i = 0; i < xw; ++i) g[i] = gp[i];
        gw = xw;
    }
    if (din) for (uint32_t i = 0; i < in_w; ++i) din[i] = g[i];
}

extern "C" void ref_mlp_backward(const uint16_t* wts, uint32_t in_w, uint32_t W, uint32_t nh, uint32_t out_pad,
                                 const uint16_t* in, uint32_t is, uint32_t n, const uint16_t* acts,
                                 const uint16_t* dL_dout, uint32_t dos, uint16_t* dL_din, uint32_t dis, double* wgrad) {
    for (uint32_t s = 0; s < n; ++s) {
        float x[128], a[8 * 128], go[16], gi[128];
        for (uint32_t i = 0; i < in_w; ++i) x[i] = h2f(in[(size_t)s * is + i]);
        for (uint32_t l = 0; l < nh; ++l) for (uint32_t o = 0; o < W; ++o) a[l * W + o] = h2f(acts[((size_t)l * n + s) * W + o]);
        for (uint32_t o = 0; o < out_pad; ++o) go[o] = h2f(dL_dout[(size_t)s * dos + o]);
        mlp_bwd_one(wts, in_w, W, nh, out_pad, x, a, go, dL_din ? gi : nullptr, wgrad);
        if (dL_din) for (uint32_t i = 0; i < in_w; ++i) dL_din[(size_t)s * dis + i] = f2h(gi[i]);
    }
}

// Xavier-uniform per layer matrix, memory order, scale sqrt(6/(fan_in+fan_out)) (tcnn initialize_xavier_uniform; recalled)
extern "C" void ref_mlp_init(uint32_t in_w, uint32_t W, uint32_t nh, uint32_t out_pad, RefPcg32* rng, float* p) {
    uint32_t cw = in_w;
    for (uint32_t l = 0; l <= nh; ++l) {
        uint32_t rows = (l == nh) ? out_pad : W;
        float s = sqrtf(6.0f / (float)(rows + cw));
        for (uint32_t i = 0; i < rows * cw; ++i) *p++ = (ref_pcg32_next_float(rng) * 2.0f - 1.0f) * s;
        cw = W;
    }
}

// ---------------------------------------------------------------------------------------------
// NerfNetwork (nerf_network.h): pos-enc -> density MLP (16 out, [0] = raw density) -> [density16 | SH16]
// -> rgb MLP (3 out, padded 16) ; output (r,g,b,sigma) (extract_density :31-43,166-172).
// Flat parameter order: density MLP, rgb MLP, pos encoding, dir encoding(=0 params) (:430-457).
// ---------------------------------------------------------------------------------------------
extern "C" uint32_t ref_nerf_pos_width(const RefNerfNetCfg* c) {
    uint32_t w = c->grid.n_levels * c->grid.n_features;
    return (w + 15) / 16 * 16;   // alignment 16 for FullyFusedMLP (nerf_network.h:93-95)
}
extern "C" uint64_t ref_nerf_n_params(const RefNerfNetCfg* c, uint64_t* od, uint64_t* orgb, uint64_t* og) {
    uint64_t nd = ref_mlp_n_params(ref_nerf_pos_width(c), c->width, c->density_hidden, 16);
    uint64_t nr = ref_mlp_n_params(32, c->width, c->rgb_hidden, 16);   // rgb input = 16 (density) + 16 (SH4) (:106)
    if (od) *od = 0;
    if (orgb) *orgb = nd;
    if (og) *og = nd + nr;
    return nd + nr + ref_grid_layout(&c->grid, nullptr, nullptr, nullptr);
}

struct NerfSample { float enc[64]; float dact[8 * 128]; float dout[16]; float rin[32]; float ract[8 * 128]; float rout[16]; };

// per-sample level cut-offs for the ref_nerf_forward / ref_nerf_backward calls that follow (set_max_level_gpu on the position encoding,
// testbed_nerf.cu:3996,4004); NULL = off.  ref_nerf_density never uses them (the occupancy-grid update runs with the cut-off cleared, :4130).
static const float* g_nerf_max_level = nullptr;
extern "C" void ref_nerf_set_max_level(const float* max_level) { g_nerf_max_level = max_level; }

static void nerf_fwd_one(const RefNerfNetCfg* c, const uint16_t* params, const float* coord, NerfSample& S, bool density_only, const float* max_level = nullptr) {
    uint64_t od, orr, og; ref_nerf_n_params(c, &od, &orr, &og);
    uint32_t pw = ref_nerf_pos_width(c);
    uint16_t ench[64];
    g_grid_max_level = max_level;
    ref_grid_forward(&c->grid, params + og, coord, 7, 1, ench, pw, pw);
    g_grid_max_level = nullptr;
    for (uint32_t i = 0; i < pw; ++i) S.enc[i] = h2f(ench[i]);
    mlp_fwd_one(params + od, pw, c->width, c->density_hidden, 16, S.enc, S.dout, S.dact);
    if (density_only) return;
    float sh[16]; sh4_one(coord + 4, sh);
    for (int i = 0; i < 16; ++i) { S.rin[i] = S.dout[i]; S.rin[16 + i] = rh(sh[i]); }
    mlp_fwd_one(params + orr, 32, c->width, c->rgb_hidden, 16, S.rin, S.rout, S.ract);
}

extern "C" void ref_nerf_forward(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n, uint16_t* out) {
    #pragma omp parallel for schedule(static)
    for (int64_t s = 0; s < (int64_t)n; ++s) {
        NerfSample S; nerf_fwd_one(c, params, coords + (size_t)s * 7, S, false, g_nerf_max_level ? g_nerf_max_level + s : nullptr);
        out[s * 4 + 0] = f2h(S.rout[0]); out[s * 4 + 1] = f2h(S.rout[1]); out[s * 4 + 2] = f2h(S.rout[2]);
        out[s * 4 + 3] = f2h(S.dout[0]);
    }
}

extern "C" void ref_nerf_density(const RefNerfNetCfg* c, const uint16_t* params, const float* pos, uint32_t ps,
                                 uint32_t n, uint16_t* out) {
    #pragma omp parallel for schedule(static)
    for (int64_t s = 0; s < (int64_t)n; ++s) {
        float coord[7] = {pos[(size_t)s * ps], pos[(size_t)s * ps + 1], pos[(size_t)s * ps + 2], 0, 0, 0, 0};
        NerfSample S; nerf_fwd_one(c, params, coord, S, true);
        out[s] = f2h(S.dout[0]);
    }
}

static void nerf_backward_impl(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n, const uint16_t* dL_dout, double* grads, float* dL_dcoords);
extern "C" void ref_nerf_backward(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n,
                                  const uint16_t* dL_dout, double* grads) { nerf_backward_impl(c, params, coords, n, dL_dout, grads, nullptr); }
// the same with NerfNetwork::backward's dL_dinput (nerf_network.h:262,282-330): dL_dcoords [n][7] fp32, position through the hash grid
// (dL/d encoding in network precision), direction through SH4 (dL/d SH in network precision); the dt column is not written
extern "C" void ref_nerf_backward_input(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n,
                                        const uint16_t* dL_dout, double* grads, float* dL_dcoords) { nerf_backward_impl(c, params, coords, n, dL_dout, grads, dL_dcoords); }
static void nerf_backward_impl(const RefNerfNetCfg* c, const uint16_t* params, const float* coords, uint32_t n,
                                  const uint16_t* dL_dout, double* grads, float* dL_dcoords) {
    uint64_t od, orr, og; ref_nerf_n_params(c, &od, &orr, &og);
    uint32_t pw = ref_nerf_pos_width(c);
    for (uint32_t s = 0; s < n; ++s) {
        NerfSample S; nerf_fwd_one(c, params, coords + (size_t)s * 7, S, false, g_nerf_max_level ? g_nerf_max_level + s : nullptr);
        // extract_rgb (nerf_network.h:45-60,271-275): only rgb channels feed the rgb net
        float g[16] = {0}; for (int i = 0; i < 3; ++i) g[i] = h2f(dL_dout[s * 4 + i]);
        float drin[32];
        mlp_bwd_one(params + orr, 32, c->width, c->rgb_hidden, 16, S.rin, S.ract, g, drin, grads + orr);
        // add_density_gradient (nerf_network.h:62-74,301-308): half + half
        float gd[16]; for (int i = 0; i < 16; ++i) gd[i] = drin[i];
        gd[0] = rh(gd[0] + h2f(dL_dout[s * 4 + 3]));
        float denc[64];
        mlp_bwd_one(params + od, pw, c->width, c->density_hidden, 16, S.enc, S.dact, gd, denc, grads + od);
        uint16_t dench[64]; for (uint32_t i = 0; i < pw; ++i) dench[i] = f2h(denc[i]);
        g_grid_max_level = g_nerf_max_level ? g_nerf_max_level + s : nullptr;
        ref_grid_backward(&c->grid, coords + (size_t)s * 7, 7, 1, dench, pw, grads + og);
        g_grid_max_level = nullptr;
        if (dL_dcoords) {
            float* dc = dL_dcoords + (size_t)s * 7;
            float J[16][3]; sh4_jacobian(coords + (size_t)s * 7 + 4, J);
            for (int a = 0; a < 3; ++a) { float acc = 0.f; for (int k = 0; k < 16; ++k) acc += rh(drin[16 + k]) * J[k][a]; dc[4 + a] = acc; }
            if (g_nerf_max_level) {   // levels the forward pass zeroed have no input derivative either
                const float ml = g_nerf_max_level[s] * (float)c->grid.n_levels + 1e-3f;
                for (uint32_t l = 0; l < c->grid.n_levels; ++l) if ((float)l >= ml) for (uint32_t f = 0; f < c->grid.n_features; ++f) dench[l * c->grid.n_features + f] = 0;
            }
            ref_grid_backward_input(&c->grid, params + og, coords + (size_t)s * 7, 7, 1, dench, pw, dc);
        }
    }
}

// initialize_params order: density MLP, rgb MLP, pos encoding (nerf_network.h:445-457); seed testbed.h:609 (1337)
extern "C" void ref_nerf_init(const RefNerfNetCfg* c, uint64_t seed, float* p) {
    uint64_t od, orr, og; ref_nerf_n_params(c, &od, &orr, &og);
    RefPcg32 r; ref_pcg32_seed(&r, seed, 1);
    ref_mlp_init(ref_nerf_pos_width(c), c->width, c->density_hidden, 16, &r, p + od);
    ref_mlp_init(32, c->width, c->rgb_hidden, 16, &r, p + orr);
    grid_init_from(&c->grid, r, p + og);
}

// ---------------------------------------------------------------------------------------------
// Optimizer: Adam with per-parameter step counts and zero-gradient skip on non-matrix (hash) params,
// L2 regularisation on matrix params only; ExponentialDecay; EMA (tcnn; call sites
// testbed_nerf.cu:3678, testbed_image.cu:283, testbed.cu:4339-4346; config configs/nerf/base.json:5-22).
// ---------------------------------------------------------------------------------------------
extern "C" void ref_adam_step(const RefAdamCfg* c, uint64_t n, float* master, uint16_t* params, const uint16_t* grads,
                              float* m, float* v, uint32_t* steps) {
    for (uint64_t i = 0; i < n; ++i) {
        float g = h2f(grads[i]) / c->loss_scale;
        if (i >= c->n_matrix_params) { if (!c->optimize_non_matrix || g == 0.0f) continue; }
        else { if (!c->optimize_matrix) continue; }
        float w = master[i];
        if (i < c->n_matrix_params) g += c->l2_reg * w;
        float g2 = g * g;
        float fm = m[i] = c->beta1 * m[i] + (1.f - c->beta1) * g;
        float sm = v[i] = c->beta2 * v[i] + (1.f - c->beta2) * g2;
        uint32_t t = ++steps[i];
        float lr = c->learning_rate * sqrtf(1.f - powf(c->beta2, (float)t)) / (1.f - powf(c->beta1, (float)t));
        float eff = lr / (sqrtf(sm) + c->epsilon);
        float nw = w - eff * fm;
        master[i] = nw;
        params[i] = f2h(nw);
    }
}

// learning-rate factor after `step` optimizer steps (step counted from 1 after the first update)
extern "C" float ref_exp_decay_lr(float base_lr, uint32_t step, uint32_t decay_start, uint32_t decay_interval,
                                  float decay_base, uint32_t decay_end) {
    float lr = base_lr;
    if (decay_interval == 0) return lr;
    for (uint32_t s = 1; s <= step; ++s)
        if (s >= decay_start && s < decay_end && (s - decay_start) % decay_interval == 0) lr *= decay_base;
    return lr;
}

extern "C" void ref_ema_step(uint64_t n, float decay, uint32_t step, const uint16_t* params, float* ema, uint16_t* ema_h) {
    float old_debias = 1.f - powf(decay, (float)(step - 1));
    float new_debias = 1.f - powf(decay, (float)step);
    for (uint64_t i = 0; i < n; ++i) {
        float f = (ema[i] * decay * old_debias + h2f(params[i]) * (1.f - decay)) / new_debias;
        ema[i] = f; ema_h[i] = f2h(f);
    }
}

// ---- OpenMP control (bench.py: torchrun exports OMP_NUM_THREADS=1, which would silently cap the CPU arm) ----
#include <omp.h>
extern "C" void ref_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }
extern "C" int ref_get_max_threads() { return omp_get_max_threads(); }
