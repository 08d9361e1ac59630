// Device helpers of the NeRF path: constants, exponential stepping, cascaded occupancy lookups, ray generation,
// colour conversions, low-discrepancy jitter.  Restates the helper layer of the reference
// (src/testbed_nerf.cu:55-184,285-315,389-457,614-633; include/.../common_device.cuh:75-121,289-581,885-925;
// random_val.cuh:151-325; bounding_box.cuh:163-223) for sm_100a; arithmetic order is kept identical to the CPU
// oracle (no implicit FMA contraction, explicit fmaf where fused) so traversal decisions are bit-exact.
#pragma once
#include "ngp_common.cuh"
#include "../../include/ngp_b200_math.h"

namespace ngp {

constexpr uint32_t NERF_GRIDSIZE = 128;
constexpr uint32_t NERF_GRID_N_CELLS = 128u * 128u * 128u;
constexpr uint32_t NERF_STEPS = 1024;
constexpr uint32_t NERF_CASCADES = 8;
constexpr uint32_t N_MAX_RANDOM_SAMPLES_PER_RAY = 16;
constexpr float NERF_MIN_OPTICAL_THICKNESS = 0.01f;
constexpr float MAX_DEPTH = 16384.0f;
constexpr float LOSS_SCALE = 128.0f;
constexpr float SQRT3 = 1.73205080757f;
NGP_HD float STEPSIZE() { return SQRT3 / (float)NERF_STEPS; }
NGP_HD float MIN_CONE_STEPSIZE() { return STEPSIZE(); }
NGP_HD float MAX_CONE_STEPSIZE() { return STEPSIZE() * (float)(1 << (NERF_CASCADES - 1)) * (float)NERF_STEPS / (float)NERF_GRIDSIZE; }

enum ImageDataType : int { IMG_BYTE = 0, IMG_FLOAT = 1 };
enum LensMode : int { LENS_PERSPECTIVE = 0, LENS_OPENCV = 1 };
enum NerfActivation : int { ACT_NONE = 0, ACT_RELU = 1, ACT_LOGISTIC = 2, ACT_EXPONENTIAL = 3 };
enum LossType : int { LOSS_L2 = 0, LOSS_L1 = 1, LOSS_MAPE = 2, LOSS_SMAPE = 3, LOSS_HUBER = 4, LOSS_LOGL1 = 5, LOSS_RELL2 = 6 };

// TrainingImageMetadata (nerf_loader.h:31-46) + TrainingXForm.start, flattened; lives in device memory
struct ImageMeta {
    const void* pixels;
    int data_type, lens_mode;
    float lens_params[4];
    int res[2];
    float principal_point[2];
    float focal_length[2];
    float xform[12];            // TrainingXForm::start (mat4x3, column major)
    float xform_end[12];        // TrainingXForm::end; equal to xform unless the data set gives transform_matrix_start / _end
    float rolling_shutter[4];   // TrainingImageMetadata::rolling_shutter (nerf_loader.h:44): pixel time = A + B u + C v + D motionblur_time
    const float* depth;         // TrainingImageMetadata::depth (nerf_loader.h:37): per-pixel depth in scene units along the view axis, or nullptr
};

// ---- stepping (what testbed_nerf.cu:124-184 computes; evaluated through struct Cone below) ----
// logf / expf come from include/ngp_b200_math.h: explicitly rounded fp32, the same source the CPU oracle compiles, so the
// cell a step falls in is the same on both sides (north_star: occupancy traversal bit-exact, cone angle included).
// t / MIN_CONE_STEPSIZE() without the generic IEEE-division sequence: q = t * (1/S) refined by one exact-residual FMA step
// is the correctly rounded quotient; verified EXHAUSTIVELY against t / S for every float in [1e-30, 2048)
// (tools/verify_fast_div.c: 1 157 627 904 values, 0 mismatches); outside that range the plain division is used.
NGP_HD float div_by_stepsize(float t) {
    constexpr float S = SQRT3 / (float)NERF_STEPS, RS = 1.0f / S;
    if (t >= 1e-30f && t < 2048.0f) { const float q = t * RS; return fmaf(fmaf(-q, S, t), RS, q); }
    return t / S;
}

// the same for either sign: round-to-nearest division and every operation of the refinement are odd in t
NGP_HD float div_by_stepsize_signed(float t) {
    constexpr float S = SQRT3 / (float)NERF_STEPS, RS = 1.0f / S;
    const float at = fabsf(t);
    if (at >= 1e-30f && at < 2048.0f) { const float q = t * RS; return fmaf(fmaf(-q, S, t), RS, q); }
    return t / S;
}

// The same three functions with everything that depends on the cone angle alone evaluated once (the reference recomputes
// log1p_c, a, b, at, bt — six logf and two expf — inside every call; per marcher step that is ~20 transcendentals on the
// serial t chain).  Same expressions on the same inputs, hence the same bits; what is left per call is one logf or one expf.
//
// The division logf(t) / log1p_c of the logarithmic regime sits on the marcher's serial chain three times per skipped voxel.  For the
// cone angle every aabb_scale > 1 scene uses (1/256, testbed_nerf.cu:3109: log1p_c = 0x3b7f8055) the quotient q = x * (1/L) refined by one
// exact-residual FMA step equals the IEEE quotient x / L for EVERY float x with 1e-30 <= |x| < 8 (tools/verify_fast_div_log1p.c, exhaustive:
// 1 723 562 816 values, 0 mismatches; logf(t) lies in [-0.84, 6.1] over the whole regime), so that form is used there — 3 dependent
// operations instead of the ~8-deep IEEE division sequence.  Any other cone angle, and x outside the verified range, divides.
constexpr uint32_t LOG1P_C_VERIFIED_BITS = 0x3b7f8055u;   // ngp_logf(1 + 1/256)
struct Cone {
    float angle, log1p_c, a, b, at, bt, rl; bool fast_div;
    NGP_HD explicit Cone(float cone_angle) : angle(cone_angle), log1p_c(0.f), a(0.f), b(0.f), at(0.f), bt(0.f), rl(0.f), fast_div(false) {
        if (cone_angle <= 1e-5f) return;
        log1p_c = ngp_logf(1.0f + cone_angle);
        rl = 1.0f / log1p_c; fast_div = ngp_f2u(log1p_c) == LOG1P_C_VERIFIED_BITS;
#ifdef NGP_AB_OLD_MATH
        fast_div = false;
#endif
        a = (ngp_logf(MIN_CONE_STEPSIZE()) - ngp_logf(log1p_c)) / log1p_c;
        b = (ngp_logf(MAX_CONE_STEPSIZE()) - ngp_logf(log1p_c)) / log1p_c;
        at = ngp_expf(a * log1p_c); bt = ngp_expf(b * log1p_c);
    }
};
NGP_HD float to_stepping_space(float t, const Cone& c) {
    if (c.angle <= 1e-5f) return div_by_stepsize(t);   // == t / MIN_CONE_STEPSIZE(), bit for bit
    if (t <= c.at) return div_by_stepsize_signed(t - c.at) + c.a;   // == (t - at) / MIN_CONE_STEPSIZE() + a
    else if (t <= c.bt) {
        const float x = ngp_logf(t), ax = fabsf(x);
        if (c.fast_div && ax >= 1e-30f && ax < 8.0f) { const float q = x * c.rl; return fmaf(fmaf(-q, c.log1p_c, x), c.rl, q); }
        return x / c.log1p_c;
    }
    else return (t - c.bt) / MAX_CONE_STEPSIZE() + c.b;
}
NGP_HD float from_stepping_space(float n, const Cone& c) {
    if (c.angle <= 1e-5f) return n * MIN_CONE_STEPSIZE();
    if (n <= c.a) return (n - c.a) * MIN_CONE_STEPSIZE() + c.at;
    else if (n <= c.b) return ngp_expf(n * c.log1p_c);
    else return (n - c.b) * MAX_CONE_STEPSIZE() + c.bt;
}
NGP_HD float advance_n_steps(float t, const Cone& c, float n) { return from_stepping_space(to_stepping_space(t, c) + n, c); }
NGP_HD float calc_dt(float t, const Cone& c) { return advance_n_steps(t, c, 1.0f) - t; }

// ---- occupancy lookup (testbed_nerf.cu:433-457,614-633) ----
// frexpf / scalbnf on the values this file feeds them, without the library's range handling: the exponent field read or
// written directly (exact for normal numbers; zero, denormals, inf and nan take the library call, so the result is the same
// for every input).
NGP_HD int float_bits(float x) {
#ifdef __CUDA_ARCH__
    return __float_as_int(x);
#else
    int i; memcpy(&i, &x, 4); return i;
#endif
}
NGP_HD float bits_float(int i) {
#ifdef __CUDA_ARCH__
    return __int_as_float(i);
#else
    float x; memcpy(&x, &i, 4); return x;
#endif
}
NGP_HD int frexp_exponent(float m) {
    const int field = (float_bits(m) >> 23) & 0xff;
    if (field != 0 && field != 0xff) return field - 126;
    int e; frexpf(m, &e); return e;
}
NGP_HD float pow2i(int e) { return bits_float((e + 127) << 23); }   // 2^e, -126 <= e <= 127

NGP_HD uint32_t mip_from_pos(V3 p, uint32_t max_cascade) {
    float m = fmaxf(fmaxf(fabsf(p.x - 0.5f), fabsf(p.y - 0.5f)), fabsf(p.z - 0.5f));
    return (uint32_t)clampi(frexp_exponent(m) + 1, 0, (int)max_cascade);
}
NGP_HD uint32_t mip_from_dt(float dt, V3 p, uint32_t max_cascade) {
    uint32_t mip = mip_from_pos(p, max_cascade);
    dt *= 2 * NERF_GRIDSIZE;
    if (dt < 1.0f) return mip;
    const int e = frexp_exponent(dt);
    // tcnn::clamp(val, lower, upper) = val < lower ? lower : (upper < val ? upper : val)   (testbed_nerf.cu:632)
    int v = (int)mip; const int lo = e, hi = (int)max_cascade;
    v = v < lo ? lo : (hi < v ? hi : v);
    return (uint32_t)v;
}
NGP_HD uint32_t cascaded_grid_idx_at(V3 p, uint32_t mip) {
    float s = pow2i(-(int)mip);   // scalbnf(1.0f, -mip), mip <= 7
    p = (p - v3(0.5f, 0.5f, 0.5f)) * s + v3(0.5f, 0.5f, 0.5f);
    int ix = (int)(p.x * (float)NERF_GRIDSIZE), iy = (int)(p.y * (float)NERF_GRIDSIZE), iz = (int)(p.z * (float)NERF_GRIDSIZE);
    if (ix < 0 || ix >= (int)NERF_GRIDSIZE || iy < 0 || iy >= (int)NERF_GRIDSIZE || iz < 0 || iz >= (int)NERF_GRIDSIZE) return 0xFFFFFFFFu;
    return morton3D((uint32_t)ix, (uint32_t)iy, (uint32_t)iz);
}
NGP_HD bool density_grid_occupied_at(V3 p, const uint8_t* __restrict__ bitfield, uint32_t mip) {
    uint32_t idx = cascaded_grid_idx_at(p, mip);
    if (idx == 0xFFFFFFFFu) return false;
    return bitfield[idx / 8 + (size_t)NERF_GRID_N_CELLS * mip / 8] & (1 << (idx % 8));
}

// ---- DDA-like skip (testbed_nerf.cu:285-315) ----
NGP_HD float distance_to_next_voxel(V3 pos, V3 dir, V3 idir, float res) {
    V3 p = res * (pos - v3(0.5f, 0.5f, 0.5f));
    float tx = (floorf(p.x + 0.5f + 0.5f * copysignf(1.0f, dir.x)) - p.x) * idir.x;
    float ty = (floorf(p.y + 0.5f + 0.5f * copysignf(1.0f, dir.y)) - p.y) * idir.y;
    float tz = (floorf(p.z + 0.5f + 0.5f * copysignf(1.0f, dir.z)) - p.z) * idir.z;
    float t = fminf(fminf(tx, ty), tz);
    return fmaxf(t / res, 0.0f);
}
NGP_HD float advance_to_next_voxel(float t, const Cone& cone, V3 pos, V3 dir, V3 idir, uint32_t mip) {
    // res = scalbnf(128, -mip) is a power of two: the division by it in distance_to_next_voxel is the exact multiplication by 2^(mip-7)
    const float res = pow2i(7 - (int)mip), inv_res = pow2i((int)mip - 7);
    V3 p = res * (pos - v3(0.5f, 0.5f, 0.5f));
    float tx = (floorf(p.x + 0.5f + 0.5f * copysignf(1.0f, dir.x)) - p.x) * idir.x;
    float ty = (floorf(p.y + 0.5f + 0.5f * copysignf(1.0f, dir.y)) - p.y) * idir.y;
    float tz = (floorf(p.z + 0.5f + 0.5f * copysignf(1.0f, dir.z)) - p.z) * idir.z;
    float t_target = t + fmaxf(fminf(fminf(tx, ty), tz) * inv_res, 0.0f);
    t = to_stepping_space(t, cone);
    t_target = to_stepping_space(t_target, cone);
    return from_stepping_space(t + ceilf(fmaxf(t_target - t, 0.5f)), cone);
}

// ---- bounding box (bounding_box.cuh:163-223) ----
NGP_HD bool aabb_contains(const Aabb& b, V3 p) {
    return p.x >= b.min[0] && p.x <= b.max[0] && p.y >= b.min[1] && p.y <= b.max[1] && p.z >= b.min[2] && p.z <= b.max[2];
}
NGP_HD void swapf(float& a, float& b) { float t = a; a = b; b = t; }
NGP_HD void aabb_ray_intersect(const Aabb& b, V3 o, V3 d, float& out_min, float& out_max) {
    const float FMAX = 3.402823466e+38f;
    float tmin = (b.min[0] - o.x) / d.x, tmax = (b.max[0] - o.x) / d.x;
    if (tmin > tmax) swapf(tmin, tmax);
    float tymin = (b.min[1] - o.y) / d.y, tymax = (b.max[1] - o.y) / d.y;
    if (tymin > tymax) swapf(tymin, tymax);
    if (tmin > tymax || tymin > tmax) { out_min = out_max = FMAX; return; }
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = (b.min[2] - o.z) / d.z, tzmax = (b.max[2] - o.z) / d.z;
    if (tzmin > tzmax) swapf(tzmin, tzmax);
    if (tmin > tzmax || tzmin > tmax) { out_min = out_max = FMAX; return; }
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    out_min = tmin; out_max = tmax;
}
NGP_HD V3 warp_position(V3 p, const Aabb& b) { return (p - ld3(b.min)) / (ld3(b.max) - ld3(b.min)); }
NGP_HD V3 unwarp_position(V3 p, const Aabb& b) { return ld3(b.min) + p * (ld3(b.max) - ld3(b.min)); }
NGP_HD V3 warp_direction(V3 d) { return (d + v3(1.f, 1.f, 1.f)) * 0.5f; }
NGP_HD float warp_dt(float dt) {
    float mx = MIN_CONE_STEPSIZE() * (float)(1 << (NERF_CASCADES - 1));
    return (dt - MIN_CONE_STEPSIZE()) / (mx - MIN_CONE_STEPSIZE());
}
NGP_HD float unwarp_dt(float dt) {
    float mx = MIN_CONE_STEPSIZE() * (float)(1 << (NERF_CASCADES - 1));
    return dt * (mx - MIN_CONE_STEPSIZE()) + MIN_CONE_STEPSIZE();
}

// ---- colour (common_device.cuh:75-121) ----
NGP_HD float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
NGP_HD float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

// ---- network output activations (testbed_nerf.cu:317-387) ----
// density -> alpha -> transmittance decide how many samples of a ray survive compaction (T < 1e-4 cut-off), an integer
// result compared bit-exact with the oracle: exp() on that chain is ngp_expf (shared source); colours use the library.
NGP_HD float alpha_from_density(float density, float dt) { return 1.f - ngp_expf(-density * dt); }
NGP_HD float logistic(float x) { return 1.0f / (1.0f + expf(-x)); }
NGP_HD float network_to_density(float v, int act) {
    switch (act) { case ACT_NONE: return v; case ACT_RELU: return v > 0.f ? v : 0.f; case ACT_LOGISTIC: return logistic(v); default: return ngp_expf(v); }
}
NGP_HD float network_to_density_derivative(float v, int act) {
    switch (act) { case ACT_NONE: return 1.f; case ACT_RELU: return v > 0.f ? 1.f : 0.f;
        case ACT_LOGISTIC: { float d = logistic(v); return d * (1 - d); } default: return expf(clampf(v, -15.f, 15.f)); }
}
NGP_HD float network_to_rgb(float v, int act) {
    switch (act) { case ACT_NONE: return v; case ACT_RELU: return v > 0.f ? v : 0.f; case ACT_LOGISTIC: return logistic(v);
        default: return expf(clampf(v, -10.f, 10.f)); }
}
NGP_HD float network_to_rgb_derivative(float v, int act) {
    switch (act) { case ACT_NONE: return 1.f; case ACT_RELU: return v > 0.f ? 1.f : 0.f;
        case ACT_LOGISTIC: { float d = logistic(v); return d * (1 - d); } default: return expf(clampf(v, -10.f, 10.f)); }
}

// ---- lens + rays (common_device.cuh:289-377,443-510,545-581) ----
NGP_HD void opencv_delta(const float* p, float u, float v, float* du, float* dv) {
    float k1 = p[0], k2 = p[1], p1 = p[2], p2 = p[3];
    float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
    float radial = k1 * r2 + k2 * r2 * r2;
    *du = u * radial + 2.f * p1 * uv + p2 * (r2 + 2.f * u2);
    *dv = v * radial + 2.f * p2 * uv + p1 * (r2 + 2.f * v2);
}
NGP_HD void opencv_undistort(const float* params, float* u, float* v) {
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
        float J00 = 1 + (dx0f[0] - dx0b[0]) / (2 * step0), J10 = (dx1f[0] - dx1b[0]) / (2 * step1);
        float J01 = (dx0f[1] - dx0b[1]) / (2 * step0), J11 = 1 + (dx1f[1] - dx1b[1]) / (2 * step1);
        float r0 = x[0] + dx[0] - x0[0], r1 = x[1] + dx[1] - x0[1];
        float inv_det = 1.0f / (J00 * J11 - J10 * J01);
        float s0 = (J11 * r0 - J10 * r1) * inv_det, s1 = (-J01 * r0 + J00 * r1) * inv_det;
        x[0] -= s0; x[1] -= s1;
        if (s0 * s0 + s1 * s1 < kMaxStepNorm) break;
    }
    *u = x[0]; *v = x[1];
}
// uv -> unnormalised ray; default args of uv_to_ray (no foveation / parallax / aperture / near plane)
// The trainable distortion map (Testbed::m_distortion, TrainableBuffer<2, 2, float>; testbed.cu:4066-4077): a res[0] x res[1] lattice of 2-D offsets
// that uv_to_ray adds to the ray direction's x, y on the image plane, looked up bilinearly at the ray's uv (Buffer2DView::at_lerp, common.h:384-400)
struct DistortionMap { const float* data; int res[2]; };
NGP_HD void distortion_at_lerp(const DistortionMap& m, float u, float v, float* out2) {
    const float xf = (float)m.res[0] * u, yf = (float)m.res[1] * v;
    const int x = (int)xf, y = (int)yf;
    const float wx = xf - (float)x, wy = yf - (float)y;
    const int x0 = clampi(x, 0, m.res[0] - 1), x1 = clampi(x + 1, 0, m.res[0] - 1), y0 = clampi(y, 0, m.res[1] - 1), y1 = clampi(y + 1, 0, m.res[1] - 1);
    const float* a = m.data + 2 * ((size_t)x0 + (size_t)y0 * m.res[0]); const float* b = m.data + 2 * ((size_t)x1 + (size_t)y0 * m.res[0]);
    const float* c = m.data + 2 * ((size_t)x0 + (size_t)y1 * m.res[0]); const float* d = m.data + 2 * ((size_t)x1 + (size_t)y1 * m.res[0]);
    for (int k = 0; k < 2; ++k) out2[k] = (1 - wx) * (1 - wy) * a[k] + wx * (1 - wy) * b[k] + (1 - wx) * wy * c[k] + wx * wy * d[k];
}
NGP_HD void uv_to_ray(const float* xform, const int* res, const float* focal, const float* principal, int lens_mode, const float* lens_params,
                      float u, float v, V3& o, V3& d, const DistortionMap* distortion = nullptr) {
    V3 dir = {(u - principal[0]) * (float)res[0] / focal[0], (v - principal[1]) * (float)res[1] / focal[1], 1.0f};
    if (lens_mode == LENS_OPENCV) opencv_undistort(lens_params, &dir.x, &dir.y);
    if (distortion && distortion->data) { float off[2]; distortion_at_lerp(*distortion, u, v, off); dir.x += off[0]; dir.y += off[1]; }   // common_device.cuh:491-493
    d = mat3_mul(xform, dir);
    o = v3(xform[9], xform[10], xform[11]);
}
// get_xform_given_rolling_shutter (common_device.cuh:401-408): the camera of ONE ray — position interpolated linearly between the start and
// end pose of the view at the ray's pixel time, rotation by quaternion slerp.  Every training ray of the reference goes through it, also when
// start == end and the shutter is (0,0,0,0): the rotation then makes the round trip matrix -> quaternion -> normalise -> matrix, which is not
// the identity in floating point, so the round trip is part of the ray.  GLM's evaluation order is restated operation by operation
// (gtc/quaternion.inl:47-72,81-123 quat_cast / mat3_cast; ext/quaternion_common.inl:41-73 slerp; ext/quaternion_geometric.inl:11-24 normalize;
// detail/type_quat.inl:17-24 dot).  q = (w, x, y, z); matrices are column major: m[c * 3 + r].
NGP_HD void quat_cast3(const float* m, float* q) {
    const float fx = m[0] - m[4] - m[8], fy = m[4] - m[0] - m[8], fz = m[8] - m[0] - m[4], fw = m[0] + m[4] + m[8];
    int biggest = 0; float fb = fw;
    if (fx > fb) { fb = fx; biggest = 1; }
    if (fy > fb) { fb = fy; biggest = 2; }
    if (fz > fb) { fb = fz; biggest = 3; }
    const float bv = sqrtf(fb + 1.0f) * 0.5f, mult = 0.25f / bv;
    // m[c][r]: m[1][2] = m[5], m[2][1] = m[7], m[2][0] = m[6], m[0][2] = m[2], m[0][1] = m[1], m[1][0] = m[3]
    if (biggest == 0) { q[0] = bv; q[1] = (m[5] - m[7]) * mult; q[2] = (m[6] - m[2]) * mult; q[3] = (m[1] - m[3]) * mult; }
    else if (biggest == 1) { q[0] = (m[5] - m[7]) * mult; q[1] = bv; q[2] = (m[1] + m[3]) * mult; q[3] = (m[6] + m[2]) * mult; }
    else if (biggest == 2) { q[0] = (m[6] - m[2]) * mult; q[1] = (m[1] + m[3]) * mult; q[2] = bv; q[3] = (m[5] + m[7]) * mult; }
    else { q[0] = (m[1] - m[3]) * mult; q[1] = (m[6] + m[2]) * mult; q[2] = (m[5] + m[7]) * mult; q[3] = bv; }
}
NGP_HD float quat_dot(const float* a, const float* b) { return (a[0] * b[0] + a[1] * b[1]) + (a[2] * b[2] + a[3] * b[3]); }
NGP_HD void xform_given_rolling_shutter(const float* start, const float* end, const float* rolling_shutter, float u, float v, float motionblur_time, float* out) {
    const float pixel_t = rolling_shutter[0] + rolling_shutter[1] * u + rolling_shutter[2] * v + rolling_shutter[3] * motionblur_time;
    float qa[4], qb[4], z[4], q[4];
    quat_cast3(start, qa); quat_cast3(end, qb);
    float cos_theta = quat_dot(qa, qb);
    for (int k = 0; k < 4; ++k) z[k] = qb[k];
    if (cos_theta < 0.0f) { for (int k = 0; k < 4; ++k) z[k] = -qb[k]; cos_theta = -cos_theta; }
    if (cos_theta > 1.0f - 1.1920928955078125e-7f) {        // glm::epsilon<float>(): linear interpolation, mix(x, y, a) = x (1 - a) + y a
        for (int k = 0; k < 4; ++k) q[k] = qa[k] * (1.0f - pixel_t) + z[k] * pixel_t;
    } else {
        const float angle = acosf(cos_theta), s0 = sinf((1.0f - pixel_t) * angle), s1 = sinf(pixel_t * angle), sd = sinf(angle);
        for (int k = 0; k < 4; ++k) q[k] = (s0 * qa[k] + s1 * z[k]) / sd;
    }
    const float len = sqrtf(quat_dot(q, q));
    if (len <= 0.0f) { q[0] = 1.0f; q[1] = q[2] = q[3] = 0.0f; }
    else { const float inv = 1.0f / len; for (int k = 0; k < 4; ++k) q[k] = q[k] * inv; }
    const float w = q[0], x = q[1], y = q[2], zz = q[3];
    const float qxx = x * x, qyy = y * y, qzz = zz * zz, qxz = x * zz, qxy = x * y, qyz = y * zz, qwx = w * x, qwy = w * y, qwz = w * zz;
    out[0] = 1.0f - 2.0f * (qyy + qzz); out[1] = 2.0f * (qxy + qwz); out[2] = 2.0f * (qxz - qwy);
    out[3] = 2.0f * (qxy - qwz); out[4] = 1.0f - 2.0f * (qxx + qzz); out[5] = 2.0f * (qyz + qwx);
    out[6] = 2.0f * (qxz + qwy); out[7] = 2.0f * (qyz - qwx); out[8] = 1.0f - 2.0f * (qxx + qyy);
    for (int k = 0; k < 3; ++k) out[9 + k] = start[9 + k] + (end[9 + k] - start[9 + k]) * pixel_t;
}
NGP_HD void pos_to_uv(const ImageMeta& im, V3 pos, float* uv) {
    const float* m = im.xform;
    V3 dir = pos - v3(m[9], m[10], m[11]);
    float a = m[0], b = m[3], c = m[6], d_ = m[1], e = m[4], f = m[7], g = m[2], h = m[5], i = m[8];
    float det = a * (e * i - f * h) - b * (d_ * i - f * g) + c * (d_ * h - e * g);
    float id = 1.0f / det;
    V3 r = {((e * i - f * h) * dir.x + (c * h - b * i) * dir.y + (b * f - c * e) * dir.z) * id,
            ((f * g - d_ * i) * dir.x + (a * i - c * g) * dir.y + (c * d_ - a * f) * dir.z) * id,
            ((d_ * h - e * g) * dir.x + (b * g - a * h) * dir.y + (a * e - b * d_) * dir.z) * id};
    r.x /= r.z; r.y /= r.z;
    float du = 0.f, dv = 0.f;
    if (im.lens_mode == LENS_OPENCV) opencv_delta(im.lens_params, r.x, r.y, &du, &dv);
    r.x += du; r.y += dv;
    uv[0] = r.x * im.focal_length[0] / (float)im.res[0] + im.principal_point[0];
    uv[1] = r.y * im.focal_length[1] / (float)im.res[1] + im.principal_point[1];
}
// read_rgba (common_device.cuh:885-925): RGBA8 is sRGB, returned linear & premultiplied; 0x00FF00FF marks masked pixels
NGP_D void read_rgba(const ImageMeta& im, float u, float v, float* rgba) {
    int px = clampi((int)(u * (float)im.res[0]), 0, im.res[0] - 1);
    int py = clampi((int)(v * (float)im.res[1]), 0, im.res[1] - 1);
    size_t idx = (size_t)px + (size_t)py * im.res[0];
    if (im.data_type == IMG_BYTE) {
        uint32_t val = __ldg(reinterpret_cast<const uint32_t*>(im.pixels) + idx);
        if (val == 0x00FF00FFu) { rgba[0] = rgba[1] = rgba[2] = rgba[3] = -1.0f; return; }
        float alpha = (float)(val >> 24) * (1.0f / 255.0f);
        rgba[0] = srgb_to_linear((float)(val & 0xFF) * (1.0f / 255.0f)) * alpha;
        rgba[1] = srgb_to_linear((float)((val >> 8) & 0xFF) * (1.0f / 255.0f)) * alpha;
        rgba[2] = srgb_to_linear((float)((val >> 16) & 0xFF) * (1.0f / 255.0f)) * alpha;
        rgba[3] = alpha;
    } else {
        float4 p = __ldg(reinterpret_cast<const float4*>(im.pixels) + idx);
        rgba[0] = p.x; rgba[1] = p.y; rgba[2] = p.z; rgba[3] = p.w;
    }
}

// read_depth (common_device.cuh:927-933): the depth map at the pixel that holds uv
NGP_D float read_depth(const ImageMeta& im, float u, float v) {
    const int px = clampi((int)(u * (float)im.res[0]), 0, im.res[0] - 1), py = clampi((int)(v * (float)im.res[1]), 0, im.res[1] - 1);
    return __ldg(im.depth + (size_t)px + (size_t)py * im.res[0]);
}

// ---- Owen-scrambled Sobol (random_val.cuh:151-291; Burley 2019) ----
NGP_HD uint32_t sobol_dim(uint32_t index, uint32_t dim) {
    uint32_t X = 0, v = 0x80000000u;
    for (uint32_t bit = 0; bit < 32; ++bit) {
        if ((index >> bit) & 1u) X ^= v;
        v = (dim == 0) ? (v >> 1) : (v ^ (v >> 1));
    }
    return X;
}
NGP_HD uint32_t hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
NGP_HD uint32_t reverse_bits32(uint32_t x) {
    x = (((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1)); x = (((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2));
    x = (((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4)); x = (((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8));
    return (x >> 16) | (x << 16);
}
NGP_HD uint32_t laine_karras(uint32_t x, uint32_t seed) {
    x += seed; x ^= x * 0x6c50b47cu; x ^= x * 0xb82f1e52u; x ^= x * 0xc7afe638u; x ^= x * 0x8d22f6e6u; return x;
}
NGP_HD uint32_t nus_base2(uint32_t x, uint32_t seed) { x = reverse_bits32(x); x = laine_karras(x, seed); return reverse_bits32(x); }
NGP_HD float ld_random_val(uint32_t index, uint32_t seed) {
    const float S = (float)(1.0 / (double)(1ull << 32));
    index = nus_base2(index, seed);
    return (float)nus_base2(sobol_dim(index, 0), hash_combine(seed, 0)) * S;
}
NGP_HD void ld_random_val_2d(uint32_t index, uint32_t seed, float* out) {
    const float S = (float)(1.0 / (double)(1ull << 32));
    index = nus_base2(index, seed);
    for (uint32_t i = 0; i < 2; ++i) out[i] = (float)nus_base2(sobol_dim(index, i), hash_combine(seed, i)) * S;
}
NGP_HD void ld_random_pixel_offset(uint32_t spp, float* off) {
    float a[2], b[2];
    ld_random_val_2d(0, 0xdeadbeefu, a); ld_random_val_2d(spp, 0xdeadbeefu, b);
    for (int i = 0; i < 2; ++i) { float o = 0.5f - a[i] + b[i]; off[i] = o - floorf(o); }
}

}  // namespace ngp
