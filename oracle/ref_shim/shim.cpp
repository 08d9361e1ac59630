// ORACLE — TEST INFRASTRUCTURE ONLY.  Compiles the reference's OWN helper code with g++ so that the CPU restatement in
// oracle/ngp_ref_*.cpp can be checked against it:
//   * include/neural-graphics-primitives/{random_val,bounding_box,common_device}.cuh are #included from the mount;
//   * the __host__/__device__ helper functions of src/testbed_nerf.cu (constants :57-100, stepping :110-184, loss functions
//     :186-283 + :1340-1355, voxel skip :285-315, activations :317-387, warps / occupancy look-ups :389-457, mips :614-633) are extracted
//     by line range into oracle/_ref/testbed_nerf_helpers.inc AT BUILD TIME (oracle/Makefile, target `ref`) — nothing of
//     the reference is stored in this repository.
// tiny-cuda-nn is absent from the mount; ref_shim/tiny-cuda-nn/*.h stub the handful of symbols these files mention
// (clamp, logistic, morton3D, pcg32, host_device_swap).  IEEE math here vs --use_fast_math in the reference's CUDA build.
#include "cuda_stub.h"
#include <cassert>
#define __expf expf
#define __logf logf
#define __powf powf
#include <neural-graphics-primitives/random_val.cuh>
#include <neural-graphics-primitives/bounding_box.cuh>
#include <neural-graphics-primitives/nerf.h>
NGP_NAMESPACE_BEGIN
#include "testbed_nerf_helpers.inc"
NGP_NAMESPACE_END

using namespace ngp;
static BoundingBox make_bb(const float* mn, const float* mx) { return BoundingBox(vec3(mn[0], mn[1], mn[2]), vec3(mx[0], mx[1], mx[2])); }

extern "C" {
// get_xform_given_rolling_shutter (common_device.cuh:401-408) through the mount's own GLM; matrices column major [12]
void shim_get_xform_given_rolling_shutter(const float* start, const float* end, const float* rs, float u, float v, float motionblur_time, float* out) {
    TrainingXForm x;
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 3; ++r) { x.start[c][r] = start[c * 3 + r]; x.end[c][r] = end[c * 3 + r]; }
    mat4x3 m = get_xform_given_rolling_shutter(x, vec4(rs[0], rs[1], rs[2], rs[3]), vec2(u, v), motionblur_time);
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 3; ++r) out[c * 3 + r] = m[c][r];
}
// rotmat(vec3) / rotvec(mat3) (common_device.cuh:698-722), what RotationAdamOptimizer composes its steps with; 3x3 column major [9]
void shim_rotmat(const float* v, float* out) {
    mat3 m = rotmat(vec3(v[0], v[1], v[2]));
    for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) out[c * 3 + r] = m[c][r];
}
void shim_rotvec(const float* in, float* out) {
    mat3 m; for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) m[c][r] = in[c * 3 + r];
    vec3 v = rotvec(m); out[0] = v.x; out[1] = v.y; out[2] = v.z;
}
float shim_ld_random_val(uint32_t index, uint32_t seed) { return ld_random_val(index, seed); }
void shim_ld_random_val_2d(uint32_t index, uint32_t seed, float* out) { vec2 v = ld_random_val_2d(index, seed); out[0] = v.x; out[1] = v.y; }
float shim_srgb_to_linear(float x) { return srgb_to_linear(x); }
float shim_linear_to_srgb(float x) { return linear_to_srgb(x); }
void shim_aabb_ray_intersect(const float* mn, const float* mx, const float* o, const float* d, float* out) {
    vec2 r = make_bb(mn, mx).ray_intersect(vec3(o[0], o[1], o[2]), vec3(d[0], d[1], d[2])); out[0] = r.x; out[1] = r.y;
}
int shim_aabb_contains(const float* mn, const float* mx, const float* p) { return make_bb(mn, mx).contains(vec3(p[0], p[1], p[2])) ? 1 : 0; }
float shim_to_stepping_space(float t, float cone) { return to_stepping_space(t, cone); }
float shim_from_stepping_space(float n, float cone) { return from_stepping_space(n, cone); }
float shim_advance_n_steps(float t, float cone, float n) { return advance_n_steps(t, cone, n); }
float shim_calc_dt(float t, float cone) { return calc_dt(t, cone); }
float shim_advance_to_next_voxel(float t, float cone, const float* pos, const float* dir, uint32_t mip) {
    vec3 d(dir[0], dir[1], dir[2]);
    return advance_to_next_voxel(t, cone, vec3(pos[0], pos[1], pos[2]), d, vec3(1.0f) / d, mip);
}
uint32_t shim_mip_from_pos(const float* p, uint32_t max_cascade) { return mip_from_pos(vec3(p[0], p[1], p[2]), max_cascade); }
uint32_t shim_mip_from_dt(float dt, const float* p, uint32_t max_cascade) { return mip_from_dt(dt, vec3(p[0], p[1], p[2]), max_cascade); }
uint32_t shim_cascaded_grid_idx_at(const float* p, uint32_t mip) { return cascaded_grid_idx_at(vec3(p[0], p[1], p[2]), mip); }
int shim_density_grid_occupied_at(const float* p, const uint8_t* bitfield, uint32_t mip) { return density_grid_occupied_at(vec3(p[0], p[1], p[2]), bitfield, mip) ? 1 : 0; }
float shim_warp_dt(float dt) { return warp_dt(dt); }
float shim_unwarp_dt(float dt) { return unwarp_dt(dt); }
float shim_network_to_rgb(float v, int act) { return network_to_rgb(v, (ENerfActivation)act); }
float shim_network_to_rgb_derivative(float v, int act) { return network_to_rgb_derivative(v, (ENerfActivation)act); }
float shim_network_to_density(float v, int act) { return network_to_density(v, (ENerfActivation)act); }
float shim_network_to_density_derivative(float v, int act) { return network_to_density_derivative(v, (ENerfActivation)act); }
// loss_and_gradient (testbed_nerf.cu:1340-1355 -> :207-283); out = loss[3], gradient[3]
void shim_loss_and_gradient(const float* target, const float* prediction, int loss_type, float* out) {
    vec3 t(target[0], target[1], target[2]), p(prediction[0], prediction[1], prediction[2]);
    LossAndGradient lg = loss_and_gradient(t, p, (ELossType)loss_type);
    out[0] = lg.loss.x; out[1] = lg.loss.y; out[2] = lg.loss.z; out[3] = lg.gradient.x; out[4] = lg.gradient.y; out[5] = lg.gradient.z;
}
}
