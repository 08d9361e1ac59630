// ORACLE — TEST INFRASTRUCTURE ONLY (see ngp_ref.h / ngp_ref_nerf.h).
// Restates /root/reference/src/testbed_nerf.cu (sampler, occupancy grid, loss/compositor, tracer) and the
// helpers it uses from common_device.cuh / random_val.cuh / bounding_box.cuh on the CPU.
// Compiled with -ffp-contract=off: every a*b+c below is two roundings unless written as fmaf(), which is
// what the CUDA kernels (compiled with --fmad=false) do as well, so occupancy traversal is bit-exact for
// cone_angle == 0 (no transcendental in the stepping) and the integer paths are bit-exact always.
#include "ngp_ref_nerf.h"
#include "../include/ngp_b200_math.h"
#include <cmath>
#include <cfloat>
#include <cstring>
#include <vector>
#include <algorithm>

// Transcendentals that feed integer decisions (cell of a marcher step, transmittance cut-off of the compositor):
// math mode 1 (default) evaluates them with the explicitly rounded fp32 logf/expf of include/ngp_b200_math.h — the same
// source the CUDA kernels compile, so GPU and oracle agree bit for bit; math mode 0 uses libm, which is what the
// reference's own helper code compiled here with g++ uses (tests/test_oracle_pins.py pins the FORMULAS in that mode).
static int g_math_mode = 1;
extern "C" void ref_set_math_mode(int shared) { g_math_mode = shared ? 1 : 0; }
extern "C" int ref_get_math_mode() { return g_math_mode; }
extern "C" float ref_shared_logf(float x) { return ngp_logf(x); }
extern "C" float ref_shared_expf(float x) { return ngp_expf(x); }
static inline float m_logf(float x) { return g_math_mode ? ngp_logf(x) : logf(x); }
static inline float m_expf(float x) { return g_math_mode ? ngp_expf(x) : expf(x); }

static inline float h2f(uint16_t h) { _Float16 v; memcpy(&v, &h, 2); return (float)v; }
static inline uint16_t f2h(float f) { _Float16 v = (_Float16)f; uint16_t h; memcpy(&h, &v, 2); return h; }

struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { return {x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
static inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
static inline float dot3(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
static inline V3 normalize3(V3 a) { float inv = 1.0f / sqrtf(dot3(a, a)); return a * inv; }
static inline float length3(V3 a) { return sqrtf(dot3(a, a)); }
static inline V3 ray_at(V3 o, V3 d, float t) { return {fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z)}; }
static inline V3 ld3(const float* p) { return {p[0], p[1], p[2]}; }
// mat3(cam) * v with cam column-major 4x3
static inline V3 mat3_mul(const float* m, V3 v) {
    return {fmaf(m[6], v.z, fmaf(m[3], v.y, m[0] * v.x)), fmaf(m[7], v.z, fmaf(m[4], v.y, m[1] * v.x)),
            fmaf(m[8], v.z, fmaf(m[5], v.y, m[2] * v.x))};
}
static inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
static inline int clampi(int v, int lo, int hi) { return std::min(std::max(v, lo), hi); }
static inline float signf_(float x) { return copysignf(1.0f, x); }

// ---- constants (testbed_nerf.cu:55-100) ----
static const float SQRT3 = 1.73205080757f;
static inline float STEPSIZE() { return SQRT3 / (float)REF_NERF_STEPS; }
static inline float MIN_CONE_STEPSIZE() { return STEPSIZE(); }
static inline float MAX_CONE_STEPSIZE() { return STEPSIZE() * (float)(1 << (REF_NERF_CASCADES - 1)) * (float)REF_NERF_STEPS / (float)REF_NERF_GRIDSIZE; }

// ---- morton (tcnn::morton3D / morton3D_invert; standard bit interleave) ----
static inline uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u; return v;
}
extern "C" uint32_t ref_morton3D(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }
extern "C" uint32_t ref_morton3D_invert(uint32_t x) {
    x = x & 0x49249249u; x = (x | (x >> 2)) & 0xc30c30c3u; x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu; x = (x | (x >> 16)) & 0x0000ffffu; return x;
}

// ---- low-discrepancy values (random_val.cuh:151-291): Owen-scrambled Sobol, Burley 2019 ----
static inline uint32_t sobol_dim(uint32_t index, uint32_t dim) {
    // dim 0: v_b = 1<<(31-b) ; dim 1: v_0 = 1<<31, v_b = v_{b-1} ^ (v_{b-1} >> 1)   (standard Sobol' directions)
    uint32_t X = 0, v = 0x80000000u;
    for (uint32_t bit = 0; bit < 32; ++bit) {
        if ((index >> bit) & 1u) X ^= v;
        v = (dim == 0) ? (v >> 1) : (v ^ (v >> 1));
    }
    return X;
}
static inline uint32_t hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
static inline uint32_t reverse_bits(uint32_t x) {
    x = (((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1)); x = (((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2));
    x = (((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4)); x = (((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8));
    return (x >> 16) | (x << 16);
}
static inline uint32_t laine_karras(uint32_t x, uint32_t seed) {
    x += seed; x ^= x * 0x6c50b47cu; x ^= x * 0xb82f1e52u; x ^= x * 0xc7afe638u; x ^= x * 0x8d22f6e6u; return x;
}
static inline uint32_t nus_base2(uint32_t x, uint32_t seed) { x = reverse_bits(x); x = laine_karras(x, seed); return reverse_bits(x); }
extern "C" float ref_ld_random_val(uint32_t index, uint32_t seed) {
    const float S = (float)(1.0 / (double)(1ull << 32));
    index = nus_base2(index, seed);
    return (float)nus_base2(sobol_dim(index, 0), hash_combine(seed, 0)) * S;
}
extern "C" void ref_ld_random_val_2d(uint32_t index, uint32_t seed, float* out) {
    const float S = (float)(1.0 / (double)(1ull << 32));
    index = nus_base2(index, seed);
    for (uint32_t i = 0; i < 2; ++i) out[i] = (float)nus_base2(sobol_dim(index, i), hash_combine(seed, i)) * S;
}
static inline void ld_random_pixel_offset(uint32_t spp, float* off) {   // random_val.cuh:318-323
    float a[2], b[2]; ref_ld_random_val_2d(0, 0xdeadbeefu, a); ref_ld_random_val_2d(spp, 0xdeadbeefu, b);
    for (int i = 0; i < 2; ++i) { float o = 0.5f - a[i] + b[i]; off[i] = o - floorf(o); }
}

// ---- colour (common_device.cuh:75-121) ----
extern "C" float ref_srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
// ---- rolling shutter: one camera per ray ----
// glm::quat_cast(mat3) -> (w, x, y, z); m is column major, m[c][r] = a[c * 3 + r]
static void quat_from_mat3(const float* a, float* q) {
    float four_x = a[0] - a[4] - a[8], four_y = a[4] - a[0] - a[8], four_z = a[8] - a[0] - a[4], four_w = a[0] + a[4] + a[8];
    int which = 0; float big = four_w;
    if (four_x > big) { big = four_x; which = 1; }
    if (four_y > big) { big = four_y; which = 2; }
    if (four_z > big) { big = four_z; which = 3; }
    float val = sqrtf(big + 1.0f) * 0.5f, mult = 0.25f / val;
    float m12 = a[1 * 3 + 2], m21 = a[2 * 3 + 1], m20 = a[2 * 3 + 0], m02 = a[0 * 3 + 2], m01 = a[0 * 3 + 1], m10 = a[1 * 3 + 0];
    switch (which) {
        case 0: q[0] = val; q[1] = (m12 - m21) * mult; q[2] = (m20 - m02) * mult; q[3] = (m01 - m10) * mult; break;
        case 1: q[0] = (m12 - m21) * mult; q[1] = val; q[2] = (m01 + m10) * mult; q[3] = (m20 + m02) * mult; break;
        case 2: q[0] = (m20 - m02) * mult; q[1] = (m01 + m10) * mult; q[2] = val; q[3] = (m12 + m21) * mult; break;
        default: q[0] = (m01 - m10) * mult; q[1] = (m20 + m02) * mult; q[2] = (m12 + m21) * mult; q[3] = val; break;
    }
}
// rotmat(vec3) / rotvec(mat3) (common_device.cuh:698-722): Rodrigues' formula from an angle-axis vector, and back through glm's quaternion
// angle() / axis() (dependencies/glm/glm/ext/quaternion_trigonometric.inl:5-29; float32: for small angles 1 - w^2 carries a relative error
// of several percent, which is part of what the reference's RotationAdamOptimizer computes).  3x3 column major.
extern "C" void ref_rotmat(const float* v, float* m) {
    float angle = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (angle == 0.0f) { for (int i = 0; i < 9; ++i) m[i] = (i % 4 == 0) ? 1.0f : 0.0f; return; }
    float x = v[0] / angle, y = v[1] / angle, z = v[2] / angle, s = sinf(angle), c = cosf(angle), oc = 1.0f - c;
    m[0] = oc * x * x + c;     m[1] = oc * x * y + z * s; m[2] = oc * z * x - y * s;
    m[3] = oc * x * y - z * s; m[4] = oc * y * y + c;     m[5] = oc * y * z + x * s;
    m[6] = oc * z * x + y * s; m[7] = oc * y * z - x * s; m[8] = oc * z * z + c;
}
extern "C" void ref_rotvec(const float* m, float* v) {
    float q[4]; quat_from_mat3(m, q);
    float angle;
    if (fabsf(q[0]) > 0.877582561890372716130286068203503191f) {
        float a = asinf(sqrtf(q[1] * q[1] + q[2] * q[2] + q[3] * q[3])) * 2.0f;
        angle = q[0] < 0.0f ? 3.14159265358979323846264338327950288f * 2.0f - a : a;
    } else angle = acosf(q[0]) * 2.0f;
    float t1 = 1.0f - q[0] * q[0];
    float ax[3] = {0.f, 0.f, 1.f};
    if (t1 > 0.0f) { float t2 = 1.0f / sqrtf(t1); ax[0] = q[1] * t2; ax[1] = q[2] * t2; ax[2] = q[3] * t2; }
    for (int k = 0; k < 3; ++k) v[k] = ax[k] * angle;
}
static float qdot(const float* a, const float* b) { float t0 = a[0] * b[0], t1 = a[1] * b[1], t2 = a[2] * b[2], t3 = a[3] * b[3]; return (t0 + t1) + (t2 + t3); }
extern "C" void ref_xform_given_rolling_shutter(const float* start, const float* end, const float* rs, float u, float v, float motionblur_time, float* out) {
    float pixel_t = rs[0] + rs[1] * u + rs[2] * v + rs[3] * motionblur_time;
    float x[4], y[4], z[4], r[4];
    quat_from_mat3(start, x); quat_from_mat3(end, y);
    memcpy(z, y, sizeof(z));
    float cos_theta = qdot(x, y);
    if (cos_theta < 0.0f) { for (int k = 0; k < 4; ++k) z[k] = -y[k]; cos_theta = -cos_theta; }
    if (cos_theta > 1.0f - FLT_EPSILON) {
        for (int k = 0; k < 4; ++k) r[k] = x[k] * (1.0f - pixel_t) + z[k] * pixel_t;       // glm::mix
    } else {
        float angle = acosf(cos_theta);
        float sa = sinf((1.0f - pixel_t) * angle), sb = sinf(pixel_t * angle), sd = sinf(angle);
        for (int k = 0; k < 4; ++k) r[k] = (sa * x[k] + sb * z[k]) / sd;
    }
    float len = sqrtf(qdot(r, r));
    if (len <= 0.0f) { r[0] = 1.0f; r[1] = r[2] = r[3] = 0.0f; }
    else { float one_over = 1.0f / len; for (int k = 0; k < 4; ++k) r[k] = r[k] * one_over; }
    float qw = r[0], qx = r[1], qy = r[2], qz = r[3];
    float qxx = qx * qx, qyy = qy * qy, qzz = qz * qz, qxz = qx * qz, qxy = qx * qy, qyz = qy * qz, qwx = qw * qx, qwy = qw * qy, qwz = qw * qz;
    out[0] = 1.0f - 2.0f * (qyy + qzz); out[1] = 2.0f * (qxy + qwz); out[2] = 2.0f * (qxz - qwy);
    out[3] = 2.0f * (qxy - qwz); out[4] = 1.0f - 2.0f * (qxx + qzz); out[5] = 2.0f * (qyz + qwx);
    out[6] = 2.0f * (qxz + qwy); out[7] = 2.0f * (qyz - qwx); out[8] = 1.0f - 2.0f * (qxx + qyy);
    for (int k = 0; k < 3; ++k) out[9 + k] = start[9 + k] + (end[9 + k] - start[9 + k]) * pixel_t;
}

static float srgb_to_linear_derivative(float s) { return s <= 0.04045f ? 1.0f / 12.92f : 2.4f / 1.055f * powf((s + 0.055f) / 1.055f, 1.4f); }   // common_device.cuh:87-93
extern "C" float ref_linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

// ---- stepping (testbed_nerf.cu:124-184) ----
extern "C" float ref_to_stepping_space(float t, float cone_angle) {
    if (cone_angle <= 1e-5f) return t / MIN_CONE_STEPSIZE();
    float log1p_c = m_logf(1.0f + cone_angle);
    float a = (m_logf(MIN_CONE_STEPSIZE()) - m_logf(log1p_c)) / log1p_c;
    float b = (m_logf(MAX_CONE_STEPSIZE()) - m_logf(log1p_c)) / log1p_c;
    float at = m_expf(a * log1p_c), bt = m_expf(b * log1p_c);
    if (t <= at) return (t - at) / MIN_CONE_STEPSIZE() + a;
    else if (t <= bt) return m_logf(t) / log1p_c;
    else return (t - bt) / MAX_CONE_STEPSIZE() + b;
}
extern "C" float ref_from_stepping_space(float n, float cone_angle) {
    if (cone_angle <= 1e-5f) return n * MIN_CONE_STEPSIZE();
    float log1p_c = m_logf(1.0f + cone_angle);
    float a = (m_logf(MIN_CONE_STEPSIZE()) - m_logf(log1p_c)) / log1p_c;
    float b = (m_logf(MAX_CONE_STEPSIZE()) - m_logf(log1p_c)) / log1p_c;
    float at = m_expf(a * log1p_c), bt = m_expf(b * log1p_c);
    if (n <= a) return (n - a) * MIN_CONE_STEPSIZE() + at;
    else if (n <= b) return m_expf(n * log1p_c);
    else return (n - b) * MAX_CONE_STEPSIZE() + bt;
}
extern "C" float ref_advance_n_steps(float t, float c, float n) { return ref_from_stepping_space(ref_to_stepping_space(t, c) + n, c); }
extern "C" float ref_calc_dt(float t, float c) { return ref_advance_n_steps(t, c, 1.0f) - t; }

// ---- occupancy lookup (testbed_nerf.cu:433-457,614-633) ----
static inline uint32_t mip_from_pos(V3 p, uint32_t max_cascade) {
    int e; float m = fmaxf(fmaxf(fabsf(p.x - 0.5f), fabsf(p.y - 0.5f)), fabsf(p.z - 0.5f));
    frexpf(m, &e);
    return (uint32_t)clampi(e + 1, 0, (int)max_cascade);
}
static inline uint32_t mip_from_dt(float dt, V3 p, uint32_t max_cascade) {
    uint32_t mip = mip_from_pos(p, max_cascade);
    dt *= 2 * REF_NERF_GRIDSIZE;
    if (dt < 1.0f) return mip;
    int e; frexpf(dt, &e);
    int v = (int)mip; const int lo = e, hi = (int)max_cascade;   // tcnn::clamp(val, lower, upper), testbed_nerf.cu:632
    v = v < lo ? lo : (hi < v ? hi : v);
    return (uint32_t)v;
}
static inline uint32_t cascaded_grid_idx_at(V3 p, uint32_t mip) {
    float s = scalbnf(1.0f, -(int)mip);
    p = (p - v3(0.5f, 0.5f, 0.5f)) * s + v3(0.5f, 0.5f, 0.5f);
    int ix = (int)(p.x * (float)REF_NERF_GRIDSIZE), iy = (int)(p.y * (float)REF_NERF_GRIDSIZE), iz = (int)(p.z * (float)REF_NERF_GRIDSIZE);
    if (ix < 0 || ix >= (int)REF_NERF_GRIDSIZE || iy < 0 || iy >= (int)REF_NERF_GRIDSIZE || iz < 0 || iz >= (int)REF_NERF_GRIDSIZE) return 0xFFFFFFFFu;
    return ref_morton3D(ix, iy, iz);
}
static inline bool occupied_at(V3 p, const uint8_t* bf, uint32_t mip) {
    uint32_t idx = cascaded_grid_idx_at(p, mip);
    if (idx == 0xFFFFFFFFu) return false;
    return bf[idx / 8 + (size_t)REF_NERF_GRID_N_CELLS * mip / 8] & (1 << (idx % 8));
}
extern "C" uint32_t ref_mip_from_pos(const float* p, uint32_t mc) { return mip_from_pos(ld3(p), mc); }
extern "C" uint32_t ref_mip_from_dt(float dt, const float* p, uint32_t mc) { return mip_from_dt(dt, ld3(p), mc); }
extern "C" uint32_t ref_cascaded_grid_idx_at(const float* p, uint32_t mip) { return cascaded_grid_idx_at(ld3(p), mip); }
extern "C" int ref_density_grid_occupied_at(const float* p, const uint8_t* bf, uint32_t mip) { return occupied_at(ld3(p), bf, mip); }

// testbed_nerf.cu:285-315
static inline float distance_to_next_voxel(V3 pos, V3 dir, V3 idir, float res) {
    V3 p = res * (pos - v3(0.5f, 0.5f, 0.5f));
    float tx = (floorf(p.x + 0.5f + 0.5f * signf_(dir.x)) - p.x) * idir.x;
    float ty = (floorf(p.y + 0.5f + 0.5f * signf_(dir.y)) - p.y) * idir.y;
    float tz = (floorf(p.z + 0.5f + 0.5f * signf_(dir.z)) - p.z) * idir.z;
    float t = fminf(fminf(tx, ty), tz);
    return fmaxf(t / res, 0.0f);
}
static inline float advance_to_next_voxel(float t, float cone, V3 pos, V3 dir, V3 idir, uint32_t mip) {
    float res = scalbnf((float)REF_NERF_GRIDSIZE, -(int)mip);
    float t_target = t + distance_to_next_voxel(pos, dir, idir, res);
    t = ref_to_stepping_space(t, cone);
    t_target = ref_to_stepping_space(t_target, cone);
    return ref_from_stepping_space(t + ceilf(fmaxf(t_target - t, 0.5f)), cone);
}

// ---- bounding box (bounding_box.cuh:163-223) ----
static inline bool aabb_contains(const RefAabb* b, V3 p) {
    return p.x >= b->min[0] && p.x <= b->max[0] && p.y >= b->min[1] && p.y <= b->max[1] && p.z >= b->min[2] && p.z <= b->max[2];
}
extern "C" void ref_aabb_ray_intersect(const RefAabb* b, const float* o, const float* d, float* out) {
    const float FMAX = 3.402823466e+38f;
    float tmin = (b->min[0] - o[0]) / d[0], tmax = (b->max[0] - o[0]) / d[0];
    if (tmin > tmax) std::swap(tmin, tmax);
    float tymin = (b->min[1] - o[1]) / d[1], tymax = (b->max[1] - o[1]) / d[1];
    if (tymin > tymax) std::swap(tymin, tymax);
    if (tmin > tymax || tymin > tmax) { out[0] = out[1] = FMAX; return; }
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = (b->min[2] - o[2]) / d[2], tzmax = (b->max[2] - o[2]) / d[2];
    if (tzmin > tzmax) std::swap(tzmin, tzmax);
    if (tmin > tzmax || tzmin > tmax) { out[0] = out[1] = FMAX; return; }
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    out[0] = tmin; out[1] = tmax;
}
static inline V3 warp_position(V3 p, const RefAabb* b) {   // aabb.relative_pos (bounding_box.cuh:86-88)
    return (p - ld3(b->min)) / (ld3(b->max) - ld3(b->min));
}
static inline V3 unwarp_position(V3 p, const RefAabb* b) { return ld3(b->min) + p * (ld3(b->max) - ld3(b->min)); }
static inline V3 warp_direction(V3 d) { return (d + v3(1.f, 1.f, 1.f)) * 0.5f; }
static inline float warp_dt(float dt) {
    float mx = MIN_CONE_STEPSIZE() * (float)(1 << (REF_NERF_CASCADES - 1));
    return (dt - MIN_CONE_STEPSIZE()) / (mx - MIN_CONE_STEPSIZE());
}
static inline float unwarp_dt(float dt) {
    float mx = MIN_CONE_STEPSIZE() * (float)(1 << (REF_NERF_CASCADES - 1));
    return dt * (mx - MIN_CONE_STEPSIZE()) + MIN_CONE_STEPSIZE();
}

// ---- lens (common_device.cuh:289-377) ----
static inline void opencv_delta(const float* p, float u, float v, float* du, float* dv) {
    float k1 = p[0], k2 = p[1], p1 = p[2], p2 = p[3];
    float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
    float radial = k1 * r2 + k2 * r2 * r2;
    *du = u * radial + 2.f * p1 * uv + p2 * (r2 + 2.f * u2);
    *dv = v * radial + 2.f * p2 * uv + p1 * (r2 + 2.f * v2);
}
static inline void opencv_undistort(const float* params, float* u, float* v) {
    const float kMaxStepNorm = 1e-10f, kRelStepSize = 1e-6f, EPS = 1.1920929e-07f;
    float x0[2] = {*u, *v}, x[2] = {*u, *v};
    for (uint32_t i = 0; i < 100; ++i) {
        float step0 = fmaxf(EPS, fabsf(kRelStepSize * x[0])), step1 = fmaxf(EPS, fabsf(kRelStepSize * x[1]));
        float dx[2], dx0b[2], dx0f[2], dx1b[2], dx1f[2];
        opencv_delta(params, x[0], x[1], &dx[0], &dx[1]);
        opencv_delta(params, x[0] - step0, x[1], &dx0b[0], &dx0b[1]);
        opencv_delta(params, x[0] + step0, x[1], &dx0f[0], &dx0f[1]);
        opencv_delta(params, x[0], x[1] - step1, &dx1b[0], &dx1b[1]);
        opencv_delta(params, x[0], x[1] + step1, &dx1f[0], &dx1f[1]);
        // J column-major as in glm: J[c][r]
        float J00 = 1 + (dx0f[0] - dx0b[0]) / (2 * step0), J10 = (dx1f[0] - dx1b[0]) / (2 * step1);
        float J01 = (dx0f[1] - dx0b[1]) / (2 * step0), J11 = 1 + (dx1f[1] - dx1b[1]) / (2 * step1);
        float r0 = x[0] + dx[0] - x0[0], r1 = x[1] + dx[1] - x0[1];
        float inv_det = 1.0f / (J00 * J11 - J10 * J01);
        // inverse(J) * r with J = [[J00 J10],[J01 J11]] (row r, col c = J[c][r])
        float s0 = (J11 * r0 - J10 * r1) * inv_det, s1 = (-J01 * r0 + J00 * r1) * inv_det;
        x[0] -= s0; x[1] -= s1;
        if (s0 * s0 + s1 * s1 < kMaxStepNorm) break;
    }
    *u = x[0]; *v = x[1];
}
// The trainable distortion map (Testbed::m_distortion; testbed.cu:4066-4077): process-wide state for the training sampler, the tracer and the camera
// gradient.  map [h][w][2] fp32 offsets added to a ray direction's image-plane x, y (common_device.cuh:491-493, Buffer2DView::at_lerp common.h:384-400);
// gradient / gradient_weight (optional) receive deposit_image_gradient's splats (common_device.cuh:124-156; testbed_nerf.cu:2088-2100)
static const float* g_distortion = nullptr; static int g_distortion_res[2] = {0, 0}; static float* g_distortion_grad = nullptr; static float* g_distortion_weight = nullptr;
extern "C" void ref_set_distortion(const float* map, int w, int h, float* gradient, float* gradient_weight) {
    g_distortion = map; g_distortion_res[0] = w; g_distortion_res[1] = h; g_distortion_grad = gradient; g_distortion_weight = gradient_weight;
}
static void distortion_at_lerp(float u, float v, float* out) {
    const int rx = g_distortion_res[0], ry = g_distortion_res[1];
    float xf = (float)rx * u, yf = (float)ry * v;
    int x = (int)xf, y = (int)yf; float wx = xf - (float)x, wy = yf - (float)y;
    auto at = [&](int px, int py, int k) { px = clampi(px, 0, rx - 1); py = clampi(py, 0, ry - 1); return g_distortion[2 * ((size_t)px + (size_t)py * rx) + k]; };
    for (int k = 0; k < 2; ++k) out[k] = (1 - wx) * (1 - wy) * at(x, y, k) + wx * (1 - wy) * at(x + 1, y, k) + (1 - wx) * wy * at(x, y + 1, k) + wx * wy * at(x + 1, y + 1, k);
}
static void uv_to_ray_impl(const RefImage* img, const float* uv, float* o, float* d, bool with_distortion);
extern "C" void ref_uv_to_ray(const RefImage* img, const float* uv, float* o, float* d) { uv_to_ray_impl(img, uv, o, d, false); }
static void uv_to_ray_impl(const RefImage* img, const float* uv, float* o, float* d, bool with_distortion) {
    V3 dir = {(uv[0] - img->principal_point[0]) * (float)img->res[0] / img->focal_length[0],
              (uv[1] - img->principal_point[1]) * (float)img->res[1] / img->focal_length[1], 1.0f};
    if (img->lens_mode == REF_LENS_OPENCV) opencv_undistort(img->lens_params, &dir.x, &dir.y);
    if (with_distortion && g_distortion) { float off[2]; distortion_at_lerp(uv[0], uv[1], off); dir.x += off[0]; dir.y += off[1]; }
    dir = mat3_mul(img->xform, dir);
    o[0] = img->xform[9]; o[1] = img->xform[10]; o[2] = img->xform[11];
    d[0] = dir.x; d[1] = dir.y; d[2] = dir.z;
}
// pos_to_uv (common_device.cuh:545-581), default args
static inline void pos_to_uv(const RefImage* img, V3 pos, float* uv) {
    const float* m = img->xform;
    V3 dir = pos - v3(m[9], m[10], m[11]);
    // inverse(mat3(cam)) * dir via cofactors (glm::inverse)
    float a = m[0], b = m[3], c = m[6], d_ = m[1], e = m[4], f = m[7], g = m[2], h = m[5], i = m[8];   // row-major view
    float det = a * (e * i - f * h) - b * (d_ * i - f * g) + c * (d_ * h - e * g);
    float id = 1.0f / det;
    V3 r = {((e * i - f * h) * dir.x + (c * h - b * i) * dir.y + (b * f - c * e) * dir.z) * id,
            ((f * g - d_ * i) * dir.x + (a * i - c * g) * dir.y + (c * d_ - a * f) * dir.z) * id,
            ((d_ * h - e * g) * dir.x + (b * g - a * h) * dir.y + (a * e - b * d_) * dir.z) * id};
    r.x /= r.z; r.y /= r.z;
    float du = 0.f, dv = 0.f;
    if (img->lens_mode == REF_LENS_OPENCV) opencv_delta(img->lens_params, r.x, r.y, &du, &dv);
    r.x += du; r.y += dv;
    uv[0] = r.x * img->focal_length[0] / (float)img->res[0] + img->principal_point[0];
    uv[1] = r.y * img->focal_length[1] / (float)img->res[1] + img->principal_point[1];
}

// read_rgba (common_device.cuh:885-925)
static inline void read_rgba(const RefImage* img, const float* uv, float* rgba) {
    int px = clampi((int)(uv[0] * (float)img->res[0]), 0, img->res[0] - 1);
    int py = clampi((int)(uv[1] * (float)img->res[1]), 0, img->res[1] - 1);
    size_t idx = (size_t)px + (size_t)py * img->res[0];
    if (img->data_type == REF_IMG_BYTE) {
        uint32_t v = ((const uint32_t*)img->pixels)[idx];
        if (v == 0x00FF00FFu) { rgba[0] = rgba[1] = rgba[2] = rgba[3] = -1.0f; return; }
        const uint8_t* b = (const uint8_t*)&v;
        float alpha = (float)b[3] * (1.0f / 255.0f);
        for (int c = 0; c < 3; ++c) rgba[c] = ref_srgb_to_linear((float)b[c] * (1.0f / 255.0f)) * alpha;
        rgba[3] = alpha;
    } else {
        const float* p = (const float*)img->pixels + idx * 4;
        for (int c = 0; c < 4; ++c) rgba[c] = p[c];
    }
}

// =============================================================================================
// Occupancy grid
// =============================================================================================
// mark_untrained_density_grid (testbed_nerf.cu:503-592)
extern "C" void ref_mark_untrained_density_grid(uint32_t n_elements, float* grid, uint32_t n_images, const RefImage* images, int clear_visible) {
    #pragma omp parallel for schedule(dynamic, 4096)
    for (int64_t ii = 0; ii < (int64_t)n_elements; ++ii) {
        uint32_t i = (uint32_t)ii;
        if (grid[i] == -1.0f) continue;
        uint32_t level = i / REF_NERF_GRID_N_CELLS, pos_idx = i % REF_NERF_GRID_N_CELLS;
        uint32_t x = ref_morton3D_invert(pos_idx >> 0), y = ref_morton3D_invert(pos_idx >> 1), z = ref_morton3D_invert(pos_idx >> 2);
        float voxel_size = scalbnf(1.0f / REF_NERF_GRIDSIZE, (int)level);
        V3 pos = (v3((float)x, (float)y, (float)z) * (1.0f / (float)REF_NERF_GRIDSIZE) - v3(0.5f, 0.5f, 0.5f)) * scalbnf(1.0f, (int)level) + v3(0.5f, 0.5f, 0.5f);
        uint32_t count = 0;
        for (uint32_t j = 0; j < n_images && count < 1; ++j) {
            const RefImage* m = &images[j];
            V3 campos = v3(m->xform[9], m->xform[10], m->xform[11]), fwd = v3(m->xform[6], m->xform[7], m->xform[8]);
            for (uint32_t k = 0; k < 8; ++k) {
                V3 corner = pos + v3((k & 1) ? voxel_size : 0.f, (k & 2) ? voxel_size : 0.f, (k & 4) ? voxel_size : 0.f);
                V3 dir = normalize3(corner - campos);
                if (dot3(dir, fwd) < 1e-4f) continue;
                float uv[2]; pos_to_uv(m, corner, uv);
                float ro[3], rd[3]; ref_uv_to_ray(m, uv, ro, rd);
                V3 rdn = normalize3(ld3(rd));
                if (uv[0] > 0.0f && uv[1] > 0.0f && uv[0] < 1.0f && uv[1] < 1.0f && length3(rdn - dir) < 1e-3f) { ++count; break; }
            }
        }
        if (clear_visible || (grid[i] < 0) != (count < 1)) grid[i] = (count >= 1) ? 1.f : -1.f;
        else grid[i] = 1.0f;
    }
}

// generate_grid_samples_nerf_nonuniform (testbed_nerf.cu:635-676)
extern "C" void ref_generate_grid_samples_nonuniform(uint32_t n_elements, uint64_t seed, uint64_t skip, uint32_t step,
                                                     const RefAabb* aabb, const float* grid_in, float* pos_out,
                                                     uint32_t* indices, uint32_t n_cascades, float thresh) {
    RefPcg32 base; ref_pcg32_seed(&base, seed, 1); ref_pcg32_advance(&base, (int64_t)skip);
    for (uint32_t i = 0; i < n_elements; ++i) {
        RefPcg32 rng = base; ref_pcg32_advance(&rng, (int64_t)((uint64_t)i * 4));
        uint32_t level = (uint32_t)(ref_pcg32_next_float(&rng) * n_cascades) % n_cascades;
        uint32_t idx = 0;
        for (int j = 0; j < 10; ++j) {
            idx = ((i + step * n_elements) * 56924617u + (uint32_t)j * 19349663u + 96925573u) % REF_NERF_GRID_N_CELLS;
            idx += level * REF_NERF_GRID_N_CELLS;
            if (grid_in[idx] > thresh) break;
        }
        uint32_t pos_idx = idx % REF_NERF_GRID_N_CELLS;
        uint32_t x = ref_morton3D_invert(pos_idx >> 0), y = ref_morton3D_invert(pos_idx >> 1), z = ref_morton3D_invert(pos_idx >> 2);
        float r0 = ref_pcg32_next_float(&rng), r1 = ref_pcg32_next_float(&rng), r2 = ref_pcg32_next_float(&rng);
        V3 pos = ((v3((float)x, (float)y, (float)z) + v3(r0, r1, r2)) * (1.0f / (float)REF_NERF_GRIDSIZE) - v3(0.5f, 0.5f, 0.5f)) * scalbnf(1.0f, (int)level) + v3(0.5f, 0.5f, 0.5f);
        V3 w = warp_position(pos, aabb);
        pos_out[(size_t)i * 3] = w.x; pos_out[(size_t)i * 3 + 1] = w.y; pos_out[(size_t)i * 3 + 2] = w.z;
        indices[i] = idx;
    }
}

static inline float logistic(float x) { return 1.0f / (1.0f + expf(-x)); }
static inline float network_to_density(float v, int act) {   // testbed_nerf.cu:354-363
    switch (act) { case REF_ACT_NONE: return v; case REF_ACT_RELU: return v > 0.f ? v : 0.f;
        case REF_ACT_LOGISTIC: return logistic(v); default: return m_expf(v); }
}
static inline float network_to_density_derivative(float v, int act) {   // :365-378
    switch (act) { case REF_ACT_NONE: return 1.f; case REF_ACT_RELU: return v > 0.f ? 1.f : 0.f;
        case REF_ACT_LOGISTIC: { float d = logistic(v); return d * (1 - d); }
        default: return expf(clampf(v, -15.f, 15.f)); }
}
static inline float network_to_rgb(float v, int act) {   // :317-327
    switch (act) { case REF_ACT_NONE: return v; case REF_ACT_RELU: return v > 0.f ? v : 0.f;
        case REF_ACT_LOGISTIC: return logistic(v); default: return expf(clampf(v, -10.f, 10.f)); }
}
static inline float network_to_rgb_derivative(float v, int act) {   // :329-343
    switch (act) { case REF_ACT_NONE: return 1.f; case REF_ACT_RELU: return v > 0.f ? 1.f : 0.f;
        case REF_ACT_LOGISTIC: { float d = logistic(v); return d * (1 - d); }
        default: return expf(clampf(v, -10.f, 10.f)); }
}

// splat_grid_samples_nerf_max_nearest_neighbor (testbed_nerf.cu:678-702): atomicMax on float bits
extern "C" void ref_splat_grid_samples(uint32_t n, const uint32_t* indices, const uint16_t* density_raw, float* grid_tmp, int dact) {
    for (uint32_t i = 0; i < n; ++i) {
        float mlp = network_to_density(h2f(density_raw[i]), dact);
        float ot = mlp * scalbnf(MIN_CONE_STEPSIZE(), 0);
        uint32_t nb, ob; memcpy(&nb, &ot, 4); memcpy(&ob, &grid_tmp[indices[i]], 4);
        if (nb > ob) memcpy(&grid_tmp[indices[i]], &nb, 4);
    }
}
// ema_grid_samples_nerf (testbed_nerf.cu:731-754)
extern "C" void ref_ema_grid_samples(uint32_t n_elements, float decay, float* grid, const float* tmp) {
    for (uint32_t i = 0; i < n_elements; ++i) { float p = grid[i]; grid[i] = (p < 0.f) ? p : fmaxf(p * decay, tmp[i]); }
}
// update_density_grid_mean_and_bitfield (testbed_nerf.cu:3538-3567) + grid_to_bitfield (:762-786) + bitfield_max_pool (:788-809)
extern "C" float ref_update_bitfield(const float* grid, uint32_t n_cascades, uint8_t* bf) {
    const uint32_t NC = REF_NERF_GRID_N_CELLS;
    double sum = 0.0;
    for (uint32_t i = 0; i < NC; ++i) sum += (double)(fmaxf(grid[i], 0.f) / (float)NC);
    float mean = (float)sum;
    float thresh = std::min(REF_NERF_MIN_OPTICAL_THICKNESS, mean);
    uint32_t n_bytes = NC / 8 * REF_NERF_CASCADES, n_nonzero = NC / 8 * n_cascades;
    for (uint32_t i = 0; i < n_bytes; ++i) {
        if (i >= n_nonzero) { bf[i] = 0; continue; }
        uint8_t bits = 0;
        for (uint32_t j = 0; j < 8; ++j) bits |= grid[(size_t)i * 8 + j] > thresh ? (uint8_t)(1u << j) : 0;
        bf[i] = bits;
    }
    for (uint32_t level = 1; level < REF_NERF_CASCADES; ++level) {
        const uint8_t* prev = bf + (size_t)NC * (level - 1) / 8;
        uint8_t* next = bf + (size_t)NC * level / 8;
        for (uint32_t i = 0; i < NC / 64; ++i) {
            uint8_t bits = 0;
            for (uint32_t j = 0; j < 8; ++j) bits |= prev[(size_t)i * 8 + j] > 0 ? (uint8_t)(1u << j) : 0;
            uint32_t x = ref_morton3D_invert(i >> 0) + REF_NERF_GRIDSIZE / 8, y = ref_morton3D_invert(i >> 1) + REF_NERF_GRIDSIZE / 8,
                     z = ref_morton3D_invert(i >> 2) + REF_NERF_GRIDSIZE / 8;
            next[ref_morton3D(x, y, z)] |= bits;
        }
    }
    return mean;
}

// =============================================================================================
// Training sampler
// =============================================================================================
struct RaySetup { bool valid; V3 o, d_unnorm, d; float startt, cone_angle; };

// shared head of generate_training_samples_nerf (testbed_nerf.cu:1417-1499) and of the loss kernel's RNG replay (:1757-1780)
// Error-map importance sampling (testbed.h:733-738; testbed_nerf.cu:1232-1338,1846,1869-1900,1969-1972,2083-2107,2354-2409): a process-wide state
// for the sampler / loss / camera-gradient functions below.  CDF pointers that are NULL leave that choice uniform; error_map NULL = no splatting.
static const float* g_cdf_x_cond_y = nullptr; static const float* g_cdf_y = nullptr; static const float* g_cdf_img = nullptr; static int g_cdf_res[2] = {0, 0};
static float* g_error_map = nullptr; static int g_error_map_res[2] = {0, 0};
extern "C" void ref_set_error_map(const float* cdf_x_cond_y, const float* cdf_y, const float* cdf_img, int cdf_w, int cdf_h, float* error_map, int map_w, int map_h) {
    g_cdf_x_cond_y = cdf_x_cond_y; g_cdf_y = cdf_y; g_cdf_img = cdf_img; g_cdf_res[0] = cdf_w; g_cdf_res[1] = cdf_h;
    g_error_map = error_map; g_error_map_res[0] = map_w; g_error_map_res[1] = map_h;
}
static uint32_t binary_search(float val, const float* data, uint32_t length) {   // common.h:268-291
    if (length == 0) return 0;
    uint32_t it, count = length, step, first = 0;
    while (count > 0) {
        it = first; step = count / 2; it += step;
        if (data[it] < val) { first = ++it; count -= step + 1; } else count = step;
    }
    return std::min(first, length - 1);
}
static inline uint32_t image_idx(uint32_t base_idx, uint32_t n_rays, uint32_t n_images, float* pdf = nullptr) {   // :1317-1338
    if (g_cdf_img) {
        float sample = ref_ld_random_val(base_idx, 0xdeadbeefu);
        uint32_t img = binary_search(sample, g_cdf_img, n_images);
        if (pdf) { float prev = img > 0 ? g_cdf_img[img - 1] : 0.0f; *pdf = (g_cdf_img[img] - prev) * (float)n_images; }
        return img;
    }
    if (pdf) *pdf = 1.0f;
    return ((base_idx * n_images) / n_rays) % n_images;   // uint32 product wraps exactly as in the reference
}
static void sample_cdf_2d(float* uv, uint32_t img, float* pdf) {   // :1232-1263
    const float UNIFORM_SAMPLING_FRACTION = 0.5f;
    if (uv[0] < UNIFORM_SAMPLING_FRACTION) { uv[0] /= UNIFORM_SAMPLING_FRACTION; return; }
    uv[0] = (uv[0] - UNIFORM_SAMPLING_FRACTION) / (1.0f - UNIFORM_SAMPLING_FRACTION);
    const uint32_t rx = (uint32_t)g_cdf_res[0], ry = (uint32_t)g_cdf_res[1];
    const float* cy = g_cdf_y + (size_t)img * ry;
    uint32_t y = binary_search(uv[1], cy, ry);
    float prev = y > 0 ? cy[y - 1] : 0.0f;
    float pmf_y = cy[y] - prev;
    uv[1] = (uv[1] - prev) / pmf_y;
    const float* cx = g_cdf_x_cond_y + (size_t)img * ry * rx + (size_t)y * rx;
    uint32_t x = binary_search(uv[0], cx, rx);
    prev = x > 0 ? cx[x - 1] : 0.0f;
    float pmf_x = cx[x] - prev;
    uv[0] = (uv[0] - prev) / pmf_x;
    if (pdf) *pdf = pmf_x * pmf_y * (float)(rx * ry);
    uv[0] = ((float)x + uv[0]) / (float)rx; uv[1] = ((float)y + uv[1]) / (float)ry;
}
static inline void random_image_pos_training(RefPcg32* rng, const int32_t* res, int snap, float* uv, uint32_t img = 0, float* pdf = nullptr) {   // :1291-1315
    uv[0] = ref_pcg32_next_float(rng); uv[1] = ref_pcg32_next_float(rng);
    if (pdf) *pdf = 1.0f;
    if (g_cdf_x_cond_y) sample_cdf_2d(uv, img, pdf);
    if (snap) {
        uv[0] = ((float)clampi((int)(uv[0] * (float)res[0]), 0, res[0] - 1) + 0.5f) / (float)res[0];
        uv[1] = ((float)clampi((int)(uv[1] * (float)res[1]), 0, res[1] - 1) + 0.5f) / (float)res[1];
    }
}

// Data-parallel sharding (no reference counterpart): rank r of G draws the global ray indices
// [r*R, (r+1)*R) of a G*R-ray step — image choice, RNG stream and loss normalisation use the GLOBAL index / count.
// (0, 0) = single process.  Mirrors ngp_testbed_set_shard of the CUDA library.
static uint32_t g_ray_offset = 0, g_n_rays_global = 0;
extern "C" void ref_set_shard(uint32_t ray_offset, uint32_t n_rays_global) { g_ray_offset = ray_offset; g_n_rays_global = n_rays_global; }
// max_level_rand_training (testbed.h:573; testbed_nerf.cu:1438,1652-1657,1775,1915-1917): when on, every ray draws a level cut-off 2u between its
// pixel position and its motion-blur time; the sampler writes it to max_level[base .. base + numsteps) and the loss kernel to
// max_level_compacted[cbase .. cbase + cn) (either pointer may be NULL)
static int g_max_level_rand = 0; static float* g_max_level = nullptr; static float* g_max_level_compacted = nullptr;
extern "C" void ref_set_max_level_rand(int on, float* max_level, float* max_level_compacted) { g_max_level_rand = on; g_max_level = max_level; g_max_level_compacted = max_level_compacted; }

extern "C" void ref_generate_training_samples(uint32_t n_rays, const RefAabb* aabb, uint32_t max_samples, uint32_t n_rays_total,
                                              uint64_t seed, uint64_t skip, uint32_t n_images, const RefImage* images,
                                              const uint8_t* bf, uint32_t max_mip, int snap, float cone_angle_constant,
                                              RefSamplerOut* out) {
    (void)n_rays_total;
    RefPcg32 base; ref_pcg32_seed(&base, seed, 1); ref_pcg32_advance(&base, (int64_t)skip);
    uint32_t ray_counter = 0, numsteps_counter = 0;
    const uint32_t n_global = g_n_rays_global ? g_n_rays_global : n_rays;
    for (uint32_t i = 0; i < n_rays; ++i) {
        const uint32_t ig = g_ray_offset + i;
        uint32_t img = image_idx(ig, n_global, n_images);
        const RefImage* m = &images[img];
        RefPcg32 rng = base; ref_pcg32_advance(&rng, (int64_t)((uint64_t)ig * REF_N_MAX_RANDOM_SAMPLES_PER_RAY));
        float uv[2]; random_image_pos_training(&rng, m->res, snap, uv, img);
        float texel[4]; read_rgba(m, uv, texel);
        if (texel[0] < 0.0f) continue;                      // masked-away pixel (:1430-1435)
        const float max_level = g_max_level_rand ? ref_pcg32_next_float(&rng) * 2.0f : 1.0f;   // :1438
        const float motionblur_time = ref_pcg32_next_float(&rng);   // :1440
        RefImage at_t = *m;                                 // the view with the camera of THIS ray (:1447-1451)
        ref_xform_given_rolling_shutter(m->xform, m->xform_end, m->rolling_shutter, uv[0], uv[1], motionblur_time, at_t.xform);
        float ro[3], rd[3]; uv_to_ray_impl(&at_t, uv, ro, rd, true);
        V3 o = ld3(ro), du = ld3(rd);
        if (du.x == 0.f && du.y == 0.f && du.z == 0.f) { o = v3(at_t.xform[9], at_t.xform[10], at_t.xform[11]); du = v3(at_t.xform[6], at_t.xform[7], at_t.xform[8]); }
        V3 d = normalize3(du);
        float tmm[2]; float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
        ref_aabb_ray_intersect(aabb, oo, dd, tmm);
        float cone = cone_angle_constant;                   // calc_cone_angle (:114-122)
        tmm[0] = fmaxf(tmm[0], 0.0f);
        float startt = ref_advance_n_steps(tmm[0], cone, ref_pcg32_next_float(&rng));
        V3 idir = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);

        // first pass: count (:1590-1612)
        uint32_t j = 0; float t = startt; V3 pos;
        while (aabb_contains(aabb, pos = ray_at(o, d, t)) && j < REF_NERF_STEPS) {
            float dt = ref_calc_dt(t, cone);
            uint32_t mip = mip_from_dt(dt, pos, max_mip);
            if (occupied_at(pos, bf, mip)) { ++j; t += dt; }
            else t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
        }
        if (j == 0) continue;
        uint32_t numsteps = j;
        uint32_t basei = numsteps_counter; numsteps_counter += numsteps;   // atomicAdd in ray order
        if (basei + numsteps > max_samples) continue;
        uint32_t ray_idx = ray_counter++;
        out->ray_indices[ray_idx] = i;
        float* r = out->rays + (size_t)ray_idx * 6; r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = du.x; r[4] = du.y; r[5] = du.z;
        out->numsteps[ray_idx * 2] = numsteps; out->numsteps[ray_idx * 2 + 1] = basei;
        // second pass: write (:1630-1650)
        t = startt; j = 0;
        V3 wd = warp_direction(d);
        while (aabb_contains(aabb, pos = ray_at(o, d, t)) && j < numsteps) {
            float dt = ref_calc_dt(t, cone);
            uint32_t mip = mip_from_dt(dt, pos, max_mip);
            if (occupied_at(pos, bf, mip)) {
                float* c = out->coords + (size_t)(basei + j) * 7;
                V3 wp = warp_position(pos, aabb);
                c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt); c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
                ++j; t += dt;
            } else t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
        }
        if (g_max_level_rand && g_max_level) for (j = 0; j < numsteps; ++j) g_max_level[basei + j] = max_level;   // :1652-1657
    }
    out->ray_counter = ray_counter; out->numsteps_counter = numsteps_counter;
}

// =============================================================================================
// Loss / compositor / gradient  (compute_loss_kernel_train_nerf, testbed_nerf.cu:1660-2012; defaults: no envmap,
// no error map, no depth supervision, no exposure, no extra dims, no max_level_rand_training)
// =============================================================================================
struct LG { V3 loss, grad; };
static inline float cs(float a, float b) { return copysignf(a, b); }
static float read_depth(const RefImage* m, const float* uv) {   // common_device.cuh:927-933 (image_pos :841-843)
    int px = std::min(std::max((int)(uv[0] * (float)m->res[0]), 0), m->res[0] - 1), py = std::min(std::max((int)(uv[1] * (float)m->res[1]), 0), m->res[1] - 1);
    return m->depth[(size_t)px + (size_t)py * m->res[0]];
}
static LG loss_and_gradient(V3 target, V3 pred, int type) {   // :207-283,1340-1355
    V3 diff = pred - target; LG r;
    float d[3] = {diff.x, diff.y, diff.z}, p[3] = {pred.x, pred.y, pred.z}, tg[3] = {target.x, target.y, target.z}, l[3], g[3];
    for (int c = 0; c < 3; ++c) {
        switch (type) {
            case REF_LOSS_RELL2: { float den = p[c] * p[c] + 1e-2f; l[c] = d[c] * d[c] / den; g[c] = 2.0f * d[c] / den; break; }
            case REF_LOSS_L1: l[c] = fabsf(d[c]); g[c] = cs(1.0f, d[c]); break;
            case REF_LOSS_MAPE: { float den = fabsf(p[c]) + 1e-2f; l[c] = fabsf(d[c]) / den; g[c] = cs(1.0f / den, d[c]); break; }
            case REF_LOSS_SMAPE: { float den = 0.5f * (fabsf(p[c]) + fabsf(tg[c])) + 1e-2f; l[c] = fabsf(d[c]) / den; g[c] = cs(1.0f / den, d[c]); break; }
            case REF_LOSS_HUBER: { const float alpha = 0.1f; float ad = fabsf(d[c]); float sq = 0.5f / alpha * d[c] * d[c];
                l[c] = (ad > alpha ? (ad - 0.5f * alpha) : sq) / 5.0f;
                g[c] = (ad > alpha ? (d[c] > 0 ? 1.0f : -1.0f) : (d[c] / alpha)) / 5.0f; break; }
            case REF_LOSS_LOGL1: { float dv = fabsf(d[c]) + 1.0f; l[c] = logf(dv); g[c] = cs(1.0f / dv, d[c]); break; }
            default: l[c] = d[c] * d[c]; g[c] = 2.0f * d[c]; break;
        }
    }
    r.loss = v3(l[0], l[1], l[2]); r.grad = v3(g[0], g[1], g[2]); return r;
}

extern "C" void ref_compute_loss(uint32_t n_rays, const RefAabb* aabb, uint32_t n_rays_total, uint64_t seed, uint64_t skip,
                                 uint32_t max_compacted, uint32_t rays_counter, const RefLossCfg* cfg,
                                 uint32_t n_images, const RefImage* images, const uint16_t* net_out,
                                 const uint32_t* ray_indices, const float* rays, const float* coords, RefLossOut* out) {
    (void)n_rays_total;
    const uint32_t n_global = g_n_rays_global ? g_n_rays_global : n_rays;
    RefPcg32 base; ref_pcg32_seed(&base, seed, 1); ref_pcg32_advance(&base, (int64_t)skip);
    uint32_t compacted_counter = 0;
    for (uint32_t i = 0; i < rays_counter; ++i) {
        uint32_t numsteps = out->numsteps[i * 2], basei = out->numsteps[i * 2 + 1];
        const float* cin = coords + (size_t)basei * 7;
        const uint16_t* no = net_out + (size_t)basei * 4;
        float T = 1.f; const float epsilon = 1e-4f;
        V3 rgb_ray = v3(0, 0, 0); V3 ray_o = ld3(rays + (size_t)i * 6);
        float depth_ray = 0.f;
        uint32_t cn = 0;
        for (; cn < numsteps; ++cn) {
            if (T < epsilon) break;
            V3 rgb = v3(network_to_rgb(h2f(no[cn * 4]), cfg->rgb_activation), network_to_rgb(h2f(no[cn * 4 + 1]), cfg->rgb_activation),
                        network_to_rgb(h2f(no[cn * 4 + 2]), cfg->rgb_activation));
            float dt = unwarp_dt(cin[cn * 7 + 3]);
            float density = network_to_density(h2f(no[cn * 4 + 3]), cfg->density_activation);
            float alpha = 1.f - m_expf(-density * dt);
            float weight = alpha * T;
            rgb_ray = rgb_ray + weight * rgb;
            depth_ray += weight * length3(unwarp_position(ld3(cin + (size_t)cn * 7), aabb) - ray_o);   // :1743,1752
            T *= (1.f - alpha);
        }
        // RNG replay (:1757-1780)
        uint32_t ray_idx = g_ray_offset + ray_indices[i];
        RefPcg32 rng = base; ref_pcg32_advance(&rng, (int64_t)((uint64_t)ray_idx * REF_N_MAX_RANDOM_SAMPLES_PER_RAY));
        float img_pdf = 1.0f, uv_pdf = 1.0f;
        uint32_t img = image_idx(ray_idx, n_global, n_images, &img_pdf);
        const RefImage* m = &images[img];
        float uv[2]; random_image_pos_training(&rng, m->res, cfg->snap_to_pixel_centers, uv, img, &uv_pdf);
        const float max_level = g_max_level_rand ? ref_pcg32_next_float(&rng) * 2.0f : 1.0f;   // :1775
        ref_pcg32_advance(&rng, 1);   // motionblur_time
        V3 bg = ld3(cfg->background_color);
        if (cfg->random_bg_color) { float a = ref_pcg32_next_float(&rng), b = ref_pcg32_next_float(&rng), c = ref_pcg32_next_float(&rng); bg = v3(a, b, c); }
        bg = v3(ref_srgb_to_linear(bg.x), ref_srgb_to_linear(bg.y), ref_srgb_to_linear(bg.z));
        float tex[4]; read_rgba(m, uv, tex);
        V3 trgb = v3(tex[0], tex[1], tex[2]); float ta = tex[3];
        // exposure_scale = exp(ln2 * exposure[img]) (:1794); NULL = the all-zero default, scale exactly 1
        V3 es = v3(1.f, 1.f, 1.f);
        if (cfg->exposure) es = v3(m_expf(0.6931471805599453f * cfg->exposure[img * 3]), m_expf(0.6931471805599453f * cfg->exposure[img * 3 + 1]), m_expf(0.6931471805599453f * cfg->exposure[img * 3 + 2]));
        V3 target;
        if (cfg->linear_colors || cfg->color_space == 0) {   // :1801-1808 (m_color_space is Linear by default, testbed.h:907)
            target = v3(es.x * trgb.x, es.y * trgb.y, es.z * trgb.z) + (1.0f - ta) * bg;
            if (!cfg->linear_colors) {
                target = v3(ref_linear_to_srgb(target.x), ref_linear_to_srgb(target.y), ref_linear_to_srgb(target.z));
                bg = v3(ref_linear_to_srgb(bg.x), ref_linear_to_srgb(bg.y), ref_linear_to_srgb(bg.z));
            }
        } else {                                              // color_space SRGB (:1809-1818)
            bg = v3(ref_linear_to_srgb(bg.x), ref_linear_to_srgb(bg.y), ref_linear_to_srgb(bg.z));
            if (ta > 0) {
                V3 q = v3(ref_linear_to_srgb(es.x * trgb.x / ta), ref_linear_to_srgb(es.y * trgb.y / ta), ref_linear_to_srgb(es.z * trgb.z / ta));
                target = q * ta + (1.0f - ta) * bg;
            } else target = bg;
        }
        if (cn == numsteps) rgb_ray = rgb_ray + T * bg;

        uint32_t cbase = compacted_counter; compacted_counter += cn;   // atomicAdd in ray order (:1830)
        cn = std::min(max_compacted - std::min(max_compacted, cbase), cn);
        out->numsteps[i * 2] = cn; out->numsteps[i * 2 + 1] = cbase;
        if (out->rgb_ray) { out->rgb_ray[i * 3] = rgb_ray.x; out->rgb_ray[i * 3 + 1] = rgb_ray.y; out->rgb_ray[i * 3 + 2] = rgb_ray.z; }
        if (cn == 0) continue;

        LG lg = loss_and_gradient(target, rgb_ray, cfg->loss_type);
        lg.loss = v3(lg.loss.x / (img_pdf * uv_pdf), lg.loss.y / (img_pdf * uv_pdf), lg.loss.z / (img_pdf * uv_pdf));   // :1846 (the gradient is NOT divided, :1858-1862)
        float mean_loss = (lg.loss.x + lg.loss.y + lg.loss.z) / 3.0f;
        if (out->loss) out->loss[i] = mean_loss / (float)n_global;
        if (g_error_map) {   // :1869-1900 (no sharpness data)
            const int ex = g_error_map_res[0], ey = g_error_map_res[1];
            float px = std::min(std::max(uv[0] * (float)ex - 0.5f, 0.0f), (float)ex - (1.0f + 1e-4f)), py = std::min(std::max(uv[1] * (float)ey - 0.5f, 0.0f), (float)ey - (1.0f + 1e-4f));
            int ix = (int)px, iy = (int)py; float wx = px - (float)ix, wy = py - (float)iy;
            int x = clampi(ix, 0, m->res[0] - 2), y = clampi(iy, 0, m->res[1] - 2);
            float* e = g_error_map + (size_t)img * ex * ey;
            e[(size_t)y * ex + x] += (1 - wx) * (1 - wy) * mean_loss; e[(size_t)y * ex + x + 1] += wx * (1 - wy) * mean_loss;
            e[(size_t)(y + 1) * ex + x] += (1 - wx) * wy * mean_loss; e[(size_t)(y + 1) * ex + x + 1] += wx * wy * mean_loss;
        }

        // depth supervision (:1848-1856): the depth map at the ray's pixel, along the unnormalised ray; views without a map give -1 = no term
        const float target_depth = length3(ld3(rays + (size_t)i * 6 + 3)) * ((cfg->depth_supervision_lambda > 0.0f && m->depth) ? read_depth(m, uv) : -1.0f);
        const LG lg_depth = loss_and_gradient(v3(target_depth, target_depth, target_depth), v3(depth_ray, depth_ray, depth_ray), cfg->depth_loss_type);
        const float depth_loss_gradient = target_depth > 0.0f ? cfg->depth_supervision_lambda * lg_depth.grad.x : 0.0f;
        float depth_ray2 = 0.f;

        float loss_scale = cfg->loss_scale / (float)n_global;
        const float output_l2_reg = cfg->rgb_activation == REF_ACT_EXPONENTIAL ? 1e-4f : 0.0f;
        const float output_l1_reg_density = cfg->mean_density < REF_NERF_MIN_OPTICAL_THICKNESS ? 1e-4f : 0.0f;

        V3 rgb_ray2 = v3(0, 0, 0); T = 1.f;
        for (uint32_t j = 0; j < cn; ++j) {
            memcpy(out->coords_compacted + (size_t)(cbase + j) * 7, cin + (size_t)j * 7, 7 * sizeof(float));
            if (g_max_level_rand && g_max_level_compacted) g_max_level_compacted[cbase + j] = max_level;   // :1915-1917
            V3 pos = unwarp_position(ld3(cin + (size_t)j * 7), aabb);
            float depth = length3(pos - ray_o);
            float dt = unwarp_dt(cin[j * 7 + 3]);
            float o0 = h2f(no[j * 4]), o1 = h2f(no[j * 4 + 1]), o2 = h2f(no[j * 4 + 2]), o3 = h2f(no[j * 4 + 3]);
            V3 rgb = v3(network_to_rgb(o0, cfg->rgb_activation), network_to_rgb(o1, cfg->rgb_activation), network_to_rgb(o2, cfg->rgb_activation));
            float density = network_to_density(o3, cfg->density_activation);
            float alpha = 1.f - m_expf(-density * dt);
            float weight = alpha * T;
            rgb_ray2 = rgb_ray2 + weight * rgb;
            depth_ray2 += weight * depth;
            T *= (1.f - alpha);
            V3 suffix = rgb_ray - rgb_ray2;
            V3 dl_drgb = weight * lg.grad;
            float g0 = loss_scale * (dl_drgb.x * network_to_rgb_derivative(o0, cfg->rgb_activation) + fmaxf(0.0f, output_l2_reg * o0));
            float g1 = loss_scale * (dl_drgb.y * network_to_rgb_derivative(o1, cfg->rgb_activation) + fmaxf(0.0f, output_l2_reg * o1));
            float g2 = loss_scale * (dl_drgb.z * network_to_rgb_derivative(o2, cfg->rgb_activation) + fmaxf(0.0f, output_l2_reg * o2));
            float dd = network_to_density_derivative(o3, cfg->density_activation);
            V3 tr = T * rgb - suffix;
            const float depth_suffix = depth_ray - depth_ray2;
            const float depth_supervision = depth_loss_gradient * (T * depth - depth_suffix);   // :1934-1935
            float dl_dmlp = dd * (dt * ((lg.grad.x * tr.x + lg.grad.y * tr.y + lg.grad.z * tr.z) + depth_supervision));
            float g3 = loss_scale * dl_dmlp + (o3 < 0.0f ? -output_l1_reg_density : 0.0f) + (o3 > -10.0f && depth < cfg->near_distance ? 1e-4f : 0.0f);
            uint16_t* go = out->dloss_doutput + (size_t)(cbase + j) * 4;
            go[0] = f2h(g0); go[1] = f2h(g1); go[2] = f2h(g2); go[3] = f2h(g3);
        }
        if (cfg->exposure_gradient) {   // :1969-1983 (assumes a symmetric loss; uv_pdf = 1 without the error-map CDFs)
            V3 dgt = v3(-lg.grad.x / uv_pdf, -lg.grad.y / uv_pdf, -lg.grad.z / uv_pdf);
            if (!cfg->linear_colors) dgt = v3(dgt.x / srgb_to_linear_derivative(target.x), dgt.y / srgb_to_linear_derivative(target.y), dgt.z / srgb_to_linear_derivative(target.z));
            cfg->exposure_gradient[img * 3 + 0] += loss_scale * dgt.x * es.x * 0.6931471805599453f;
            cfg->exposure_gradient[img * 3 + 1] += loss_scale * dgt.y * es.y * 0.6931471805599453f;
            cfg->exposure_gradient[img * 3 + 2] += loss_scale * dgt.z * es.z * 0.6931471805599453f;
        }
    }
    out->compacted_counter = compacted_counter;
}

// compute_cam_gradient_train_nerf (testbed_nerf.cu:2014-2123), extrinsics part (no distortion map, no focal-length gradient, uv_pdf = 1):
// numsteps are the compacted (count, base) pairs, coords / coords_gradient the compacted batch [*][7]
static uint64_t g_cam_seed = 0, g_cam_skip = 0; static const RefImage* g_cam_images = nullptr; static int g_cam_snap = 0;
// the rng stream and views of the step, for the pixel pdf of each ray when the error-map CDFs are in use (:2083-2086)
extern "C" void ref_set_cam_gradient_rng(uint64_t seed, uint64_t skip, const RefImage* images, int snap) { g_cam_seed = seed; g_cam_skip = skip; g_cam_images = images; g_cam_snap = snap; }
extern "C" void ref_compute_cam_gradient(uint32_t n_rays, const RefAabb* aabb, uint32_t n_images, uint32_t rays_counter, const uint32_t* ray_indices,
                                         const float* rays, const uint32_t* numsteps, const float* coords, const float* coords_gradient,
                                         float* cam_pos_gradient, float* cam_rot_gradient) {
    const uint32_t n_global = g_n_rays_global ? g_n_rays_global : n_rays;
    for (uint32_t i = 0; i < rays_counter; ++i) {
        const uint32_t n = numsteps[i * 2], base = numsteps[i * 2 + 1];
        if (n == 0) continue;
        const uint32_t img = image_idx(g_ray_offset + ray_indices[i], n_global, n_images);
        float uv_pdf = 1.0f; float uv[2] = {0.f, 0.f};
        if ((g_cdf_x_cond_y || g_distortion_grad) && g_cam_images) {
            RefPcg32 rng; ref_pcg32_seed(&rng, g_cam_seed, 1); ref_pcg32_advance(&rng, (int64_t)g_cam_skip);
            ref_pcg32_advance(&rng, (int64_t)((uint64_t)(g_ray_offset + ray_indices[i]) * REF_N_MAX_RANDOM_SAMPLES_PER_RAY));
            random_image_pos_training(&rng, g_cam_images[img].res, g_cam_snap, uv, img, &uv_pdf);
        }
        const V3 ray_o = ld3(rays + (size_t)i * 6), ray_d = normalize3(ld3(rays + (size_t)i * 6 + 3));
        V3 go = v3(0, 0, 0), gd = v3(0, 0, 0);
        for (uint32_t j = 0; j < n; ++j) {
            const float* c = coords + (size_t)(base + j) * 7; const float* g = coords_gradient + (size_t)(base + j) * 7;
            const V3 diag = v3(aabb->max[0] - aabb->min[0], aabb->max[1] - aabb->min[1], aabb->max[2] - aabb->min[2]);
            const V3 pos_gradient = v3(g[0] * (1.0f / diag.x), g[1] * (1.0f / diag.y), g[2] * (1.0f / diag.z));   // warp_position_derivative
            go = go + pos_gradient;
            const V3 pos = unwarp_position(ld3(c), aabb);
            const float t = length3(pos - ray_o);
            const V3 dir_gradient = v3(g[4] * 0.5f, g[5] * 0.5f, g[6] * 0.5f);                                       // warp_direction_derivative
            gd = gd + t * pos_gradient + dir_gradient;
        }
        if (g_distortion_grad && g_cam_images) {   // :2088-2100
            float dp = gd.x * ray_d.x + gd.y * ray_d.y + gd.z * ray_d.z;
            V3 og = gd - dp * ray_d;
            const float* m = g_cam_images[img].xform;
            float a = m[0], b = m[3], c = m[6], d_ = m[1], e = m[4], f = m[7], g = m[2], h = m[5], k = m[8];
            float id = 1.0f / (a * (e * k - f * h) - b * (d_ * k - f * g) + c * (d_ * h - e * g));
            float ix = ((e * k - f * h) * og.x + (c * h - b * k) * og.y + (b * f - c * e) * og.z) * id;
            float iy = ((f * g - d_ * k) * og.x + (a * k - c * g) * og.y + (c * d_ - a * f) * og.z) * id;
            float val[2] = {ix / uv_pdf, iy / uv_pdf};
            const int rx = g_distortion_res[0], ry = g_distortion_res[1];
            float xf = (float)rx * uv[0], yf = (float)ry * uv[1];
            int tx = (int)xf, ty = (int)yf; float wx = xf - (float)tx, wy = yf - (float)ty;
            float w4[4] = {(1 - wx) * (1 - wy), wx * (1 - wy), (1 - wx) * wy, wx * wy};
            for (int q = 0; q < 4; ++q) {
                int px = clampi(tx + (q & 1), 0, rx - 1), py = clampi(ty + (q >> 1), 0, ry - 1);
                size_t o = 2 * ((size_t)px + (size_t)py * rx);
                for (int cc = 0; cc < 2; ++cc) { g_distortion_grad[o + cc] += val[cc] * w4[q]; g_distortion_weight[o + cc] += w4[q]; }
            }
        }
        if (cam_pos_gradient) { cam_pos_gradient[img * 3] += go.x / uv_pdf; cam_pos_gradient[img * 3 + 1] += go.y / uv_pdf; cam_pos_gradient[img * 3 + 2] += go.z / uv_pdf; }
        if (cam_rot_gradient) {
            cam_rot_gradient[img * 3] += (ray_d.y * gd.z - ray_d.z * gd.y) / uv_pdf;
            cam_rot_gradient[img * 3 + 1] += (ray_d.z * gd.x - ray_d.x * gd.z) / uv_pdf;
            cam_rot_gradient[img * 3 + 2] += (ray_d.x * gd.y - ray_d.y * gd.x) / uv_pdf;
        }
    }
}

// construct_cdf_2d + construct_cdf_1d (testbed_nerf.cu:2354-2409) and the host-side view CDF (:3729-3743)
extern "C" void ref_construct_error_cdfs(uint32_t n_images, uint32_t height, uint32_t width, const float* data, float* cdf_x_cond_y, float* cdf_y, float* cdf_img) {
    const float MIN_PDF = 0.01f, MIN_PMF = 0.1f;
    for (uint32_t img = 0; img < n_images; ++img) {
        for (uint32_t y = 0; y < height; ++y) {
            const size_t off = ((size_t)img * height + y) * width;
            float cum = 0;
            for (uint32_t x = 0; x < width; ++x) { cum += data[off + x] + 1e-10f; cdf_x_cond_y[off + x] = cum; }
            cdf_y[(size_t)img * height + y] = cum;
            float norm = 1.0f / cum;
            for (uint32_t x = 0; x < width; ++x) cdf_x_cond_y[off + x] = (1.0f - MIN_PDF) * cdf_x_cond_y[off + x] * norm + MIN_PDF * (float)(x + 1) / (float)width;
        }
        float* c = cdf_y + (size_t)img * height;
        float cum = 0;
        for (uint32_t y = 0; y < height; ++y) { cum += c[y]; c[y] = cum; }
        cdf_img[img] = cum;
        float norm = 1.0f / cum;
        for (uint32_t y = 0; y < height; ++y) c[y] = (1.0f - MIN_PDF) * c[y] * norm + MIN_PDF * (float)(y + 1) / (float)height;
    }
    float cum = 0;
    std::vector<float> c(n_images);
    for (uint32_t i = 0; i < n_images; ++i) { cum += cdf_img[i]; c[i] = cum; }
    float norm = 1.0f / cum;
    for (uint32_t i = 0; i < n_images; ++i) cdf_img[i] = (1.0f - MIN_PMF) * c[i] * norm + MIN_PMF * (float)(i + 1) / (float)n_images;
}

// fill_rollover_and_rescale / fill_rollover (tcnn common_device; recalled — unpinned): entries [n_input, n_target)
// are cyclic copies of [0, n_input); the rescale variant multiplies the COPIES by n_input/n_target.
extern "C" void ref_fill_rollover_half(uint32_t n_target, uint32_t stride, uint32_t n_input, uint16_t* io, int rescale) {
    size_t ni = (size_t)n_input * stride, nt = (size_t)n_target * stride;
    if (ni == 0) return;
    for (size_t i = ni; i < nt; ++i) {
        float v = h2f(io[i % ni]);
        if (rescale) v = v * (float)ni / (float)nt;
        io[i] = f2h(v);
    }
}
extern "C" void ref_fill_rollover_float(uint32_t n_target, uint32_t stride, uint32_t n_input, float* io) {
    size_t ni = (size_t)n_input * stride, nt = (size_t)n_target * stride;
    if (ni == 0) return;
    for (size_t i = ni; i < nt; ++i) io[i] = io[i % ni];
}
// testbed_nerf.cu:3605-3606
extern "C" uint32_t ref_next_rays_per_batch(uint32_t rpb, uint32_t target, uint32_t measured) {
    uint32_t r = (uint32_t)((float)rpb * (float)target / (float)measured);
    r = (r + REF_BATCH_GRANULARITY - 1) / REF_BATCH_GRANULARITY * REF_BATCH_GRANULARITY;
    return std::min(r, 1u << 18);
}

// =============================================================================================
// Inference render: init_rays_with_payload_kernel_nerf (:2228-2352) -> advance_pos_nerf (:850-893) ->
// [generate_next_nerf_network_inputs (:948-1022) -> network -> composite_kernel_nerf (:1024-1230)]* -> shade_kernel_nerf (:2164-2196)
// Per ray the wavefront order does not change the result, so the oracle walks each ray to completion.
// =============================================================================================
static float advance_to_occupied(float t, float cone, V3 o, V3 d, V3 idir, const uint8_t* bf, uint32_t min_mip, uint32_t max_mip, const RefAabb* aabb) {
    while (true) {   // if_unoccupied_advance_to_next_occupied_voxel (:811-845), MIP_FROM_DT = false
        V3 pos = ray_at(o, d, t);
        if (t >= REF_MAX_DEPTH || !aabb_contains(aabb, pos)) return REF_MAX_DEPTH;
        uint32_t mip = (uint32_t)clampi((int)mip_from_pos(pos, REF_NERF_CASCADES - 1), (int)min_mip, (int)max_mip);
        if (!bf || occupied_at(pos, bf, mip)) return t;
        while (mip < max_mip && !occupied_at(pos, bf, mip + 1)) ++mip;
        t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
    }
}

extern "C" void ref_render_nerf(const RefRenderCfg* c, const RefNerfNetCfg* net, const uint16_t* params, const uint8_t* bf,
                                float* rgba_out, float* depth_out) {
    const int W = c->res[0], H = c->res[1];
    #pragma omp parallel for schedule(dynamic, 64)
    for (int64_t pi = (int64_t)c->y0 * W; pi < (int64_t)c->y1 * W; ++pi) {
        uint32_t idx = (uint32_t)pi; int x = idx % W, y = idx / W;
        float* fb = rgba_out + (size_t)idx * 4;
        fb[0] = fb[1] = fb[2] = fb[3] = 0.f; depth_out[idx] = REF_MAX_DEPTH;
        float off[2]; ld_random_pixel_offset(c->snap_to_pixel_centers ? 0 : c->spp_index, off);
        float uv[2] = {((float)x + off[0]) / (float)W, ((float)y + off[1]) / (float)H};
        RefImage cam; memset(&cam, 0, sizeof(cam));
        cam.lens_mode = REF_LENS_PERSPECTIVE; cam.res[0] = W; cam.res[1] = H;
        cam.principal_point[0] = c->screen_center[0]; cam.principal_point[1] = c->screen_center[1];
        cam.focal_length[0] = c->focal_length[0]; cam.focal_length[1] = c->focal_length[1];
        memcpy(cam.xform, c->camera, sizeof(cam.xform));
        float ro[3], rd[3]; uv_to_ray_impl(&cam, uv, ro, rd, true);   // (render_with_lens_distortion: the caller installs the map or not)
        V3 o = ld3(ro), d = normalize3(ld3(rd));
        float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z}, tmm[2];
        ref_aabb_ray_intersect(&c->render_aabb, oo, dd, tmm);
        float t = fmaxf(tmm[0], 0.0f) + 1e-6f;
        if (!aabb_contains(&c->render_aabb, ray_at(o, d, t))) continue;
        V3 idir = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
        float cone = c->cone_angle_constant;
        // advance_pos_nerf: jittered first step + skip to first occupied voxel
        t = ref_advance_n_steps(t, cone, ref_ld_random_val(c->spp_index, idx * 786433u));
        t = advance_to_occupied(t, cone, o, d, idir, bf, 0, c->max_mip, &c->render_aabb);
        if (t >= REF_MAX_DEPTH) continue;
        float r = 0, g = 0, b = 0, a = 0, max_weight = 0, ldepth = 0; bool done = false;
        V3 cam_fwd = v3(c->camera[6], c->camera[7], c->camera[8]), cam_pos = v3(c->camera[9], c->camera[10], c->camera[11]);
        for (uint32_t step = 0; step < 10000 && !done; ++step) {
            t = advance_to_occupied(t, cone, o, d, idir, bf, 0, c->max_mip, &c->render_aabb);
            if (t >= REF_MAX_DEPTH) break;
            float dt = ref_calc_dt(t, cone);
            V3 wp = warp_position(ray_at(o, d, t), &c->train_aabb);
            V3 wd = warp_direction(d);
            float coord[7] = {wp.x, wp.y, wp.z, warp_dt(dt), wd.x, wd.y, wd.z};
            uint16_t outh[4]; ref_nerf_forward(net, params, coord, 1, outh);
            t += dt;
            // composite_kernel_nerf
            V3 pos = unwarp_position(wp, &c->train_aabb);
            float T = 1.f - a;
            float dtu = unwarp_dt(coord[3]);
            float alpha = 1.f - m_expf(-network_to_density(h2f(outh[3]), c->density_activation) * dtu);
            float weight = alpha * T;
            r += network_to_rgb(h2f(outh[0]), c->rgb_activation) * weight;
            g += network_to_rgb(h2f(outh[1]), c->rgb_activation) * weight;
            b += network_to_rgb(h2f(outh[2]), c->rgb_activation) * weight;
            a += weight;
            if (weight > max_weight) { max_weight = weight; ldepth = dot3(cam_fwd, pos - cam_pos); }
            if (a > (1.0f - c->min_transmittance)) { r /= a; g /= a; b /= a; a /= a; done = true; }
        }
        if (!(a > 0.001f)) continue;   // compact_kernel_nerf keeps only rays with alpha > 0.001 (:2220)
        if (!c->linear_colors) { r = ref_srgb_to_linear(r); g = ref_srgb_to_linear(g); b = ref_srgb_to_linear(b); }   // shade_kernel_nerf
        fb[0] = r; fb[1] = g; fb[2] = b; fb[3] = a;
        if (a > 0.2f) depth_out[idx] = ldepth;
    }
}

// ---- exports of the scalar helpers for the reference-pinning tests (tests/test_oracle_pins.py, tests/golden/) ----
extern "C" float ref_advance_to_next_voxel(float t, float cone, const float* pos, const float* dir, uint32_t mip) {
    V3 d = ld3(dir);
    return advance_to_next_voxel(t, cone, ld3(pos), d, v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z), mip);
}
extern "C" float ref_warp_dt(float dt) { return warp_dt(dt); }
extern "C" float ref_unwarp_dt(float dt) { return unwarp_dt(dt); }
extern "C" float ref_network_to_rgb(float v, int act) { return network_to_rgb(v, act); }
extern "C" float ref_network_to_rgb_derivative(float v, int act) { return network_to_rgb_derivative(v, act); }
extern "C" float ref_network_to_density(float v, int act) { return network_to_density(v, act); }
extern "C" float ref_network_to_density_derivative(float v, int act) { return network_to_density_derivative(v, act); }
extern "C" void ref_loss_and_gradient(const float* target, const float* prediction, int loss_type, float* out) {
    LG lg = loss_and_gradient(ld3(target), ld3(prediction), loss_type);
    out[0] = lg.loss.x; out[1] = lg.loss.y; out[2] = lg.loss.z; out[3] = lg.grad.x; out[4] = lg.grad.y; out[5] = lg.grad.z;
}
extern "C" int ref_aabb_contains(const RefAabb* b, const float* p) { return aabb_contains(b, ld3(p)) ? 1 : 0; }
