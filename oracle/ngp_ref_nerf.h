// ORACLE — TEST INFRASTRUCTURE ONLY (see ngp_ref.h).
// NeRF sampler / occupancy grid / loss+compositor / inference tracer restated from
// /root/reference/src/testbed_nerf.cu (in-mount; line citations at each function in ngp_ref_nerf.cpp).
#pragma once
#include "ngp_ref.h"

#ifdef __cplusplus
extern "C" {
#endif

#define REF_NERF_GRIDSIZE 128u                 // nerf.h:24-26
#define REF_NERF_GRID_N_CELLS (128u*128u*128u)
#define REF_NERF_STEPS 1024u                   // testbed_nerf.cu:58
#define REF_NERF_CASCADES 8u                   // testbed_nerf.cu:59
#define REF_N_MAX_RANDOM_SAMPLES_PER_RAY 16u   // testbed_nerf.cu:86-88
#define REF_NERF_MIN_OPTICAL_THICKNESS 0.01f   // testbed_nerf.cu:93-95
#define REF_MAX_DEPTH 16384.0f                 // common_device.cuh:33
#define REF_LOSS_SCALE 128.0f                  // testbed.h:264

enum { REF_IMG_BYTE = 0, REF_IMG_FLOAT = 1 };          // EImageDataType subset
enum { REF_LENS_PERSPECTIVE = 0, REF_LENS_OPENCV = 1 }; // ELensMode subset (common.h:243-250)
enum { REF_ACT_NONE = 0, REF_ACT_RELU = 1, REF_ACT_LOGISTIC = 2, REF_ACT_EXPONENTIAL = 3 };  // ENerfActivation
enum { REF_LOSS_L2 = 0, REF_LOSS_L1 = 1, REF_LOSS_MAPE = 2, REF_LOSS_SMAPE = 3, REF_LOSS_HUBER = 4,
       REF_LOSS_LOGL1 = 5, REF_LOSS_RELL2 = 6 };       // ELossType

// TrainingImageMetadata (nerf_loader.h:31-46) + TrainingXForm.start (common.h:232-239), flattened
typedef struct RefImage {
    const void* pixels;        // RGBA8 or RGBA float32, row-major
    int32_t data_type;         // REF_IMG_*
    int32_t lens_mode;         // REF_LENS_*
    float lens_params[4];      // k1,k2,p1,p2
    int32_t res[2];
    float principal_point[2];
    float focal_length[2];
    float xform[12];           // mat4x3 column-major: x-axis, y-axis, z-axis, position (TrainingXForm::start)
    float xform_end[12];       // TrainingXForm::end (== xform unless the data set has transform_matrix_end)
    float rolling_shutter[4];  // TrainingImageMetadata::rolling_shutter (nerf_loader.h:44)
    const float* depth;        // TrainingImageMetadata::depth (nerf_loader.h:37): [h][w] depth in scene units, or NULL
} RefImage;
// get_xform_given_rolling_shutter (common_device.cuh:401-408) with GLM's quat_cast / slerp / normalize / mat3_cast restated
// (dependencies/glm/glm/gtc/quaternion.inl:47-72,81-123; ext/quaternion_common.inl:41-73; ext/quaternion_geometric.inl:11-24)
// rotmat(vec3) / rotvec(mat3) (common_device.cuh:698-722; glm angle / axis): angle-axis vector <-> 3x3 rotation (column major)
void ref_rotmat(const float* v3, float* m9);
void ref_rotvec(const float* m9, float* v3);
void ref_xform_given_rolling_shutter(const float* start, const float* end, const float* rolling_shutter, float u, float v, float motionblur_time, float* out12);

typedef struct RefAabb { float min[3]; float max[3]; } RefAabb;

uint32_t ref_morton3D(uint32_t x, uint32_t y, uint32_t z);
uint32_t ref_morton3D_invert(uint32_t x);

// stepping helpers (testbed_nerf.cu:124-184,285-315)
float ref_to_stepping_space(float t, float cone_angle);
float ref_from_stepping_space(float n, float cone_angle);
float ref_advance_n_steps(float t, float cone_angle, float n);
float ref_calc_dt(float t, float cone_angle);
uint32_t ref_mip_from_pos(const float* pos, uint32_t max_cascade);
uint32_t ref_mip_from_dt(float dt, const float* pos, uint32_t max_cascade);
uint32_t ref_cascaded_grid_idx_at(const float* pos, uint32_t mip);
int ref_density_grid_occupied_at(const float* pos, const uint8_t* bitfield, uint32_t mip);
void ref_aabb_ray_intersect(const RefAabb* aabb, const float* o, const float* d, float* tminmax);
float ref_ld_random_val(uint32_t index, uint32_t seed);
void ref_ld_random_val_2d(uint32_t index, uint32_t seed, float* out);
float ref_srgb_to_linear(float x);
float ref_linear_to_srgb(float x);
// uv -> unnormalised ray (o[3], d[3]) (common_device.cuh:443-510, default args)
void ref_uv_to_ray(const RefImage* img, const float* uv, float* o, float* d);

// --- occupancy grid (testbed_nerf.cu:503-592,635-809,3412-3567) ---
void ref_mark_untrained_density_grid(uint32_t n_elements, float* grid, uint32_t n_images, const RefImage* images,
                                     int clear_visible_voxels);
void ref_generate_grid_samples_nonuniform(uint32_t n_elements, uint64_t rng_seed, uint64_t rng_skip, uint32_t step,
                                          const RefAabb* aabb, const float* grid_in, float* pos_out /*[n][3]*/,
                                          uint32_t* indices, uint32_t n_cascades, float thresh);
void ref_splat_grid_samples(uint32_t n, const uint32_t* indices, const uint16_t* density_raw, float* grid_tmp,
                            int density_activation);
void ref_ema_grid_samples(uint32_t n_elements, float decay, float* grid, const float* grid_tmp);
// mean over cascade 0, bitfield + 7 max-pooled mips; bitfield size REF_NERF_GRID_N_CELLS*8/8 bytes
float ref_update_bitfield(const float* grid, uint32_t n_cascades, uint8_t* bitfield);

// --- training sampler (testbed_nerf.cu:1291-1338,1382-1658), deterministic ray-order compaction ---
typedef struct RefSamplerOut {
    uint32_t* ray_indices;   // [n_rays]
    float* rays;             // [n_rays][6] unnormalised o,d
    uint32_t* numsteps;      // [n_rays][2] (numsteps, base)
    float* coords;           // [max_samples][7]
    uint32_t ray_counter;
    uint32_t numsteps_counter;   // total steps of all rays with >0 steps (may exceed max_samples)
} RefSamplerOut;
void ref_generate_training_samples(uint32_t n_rays, const RefAabb* aabb, uint32_t max_samples, uint32_t n_rays_total,
                                   uint64_t rng_seed, uint64_t rng_skip, uint32_t n_images, const RefImage* images,
                                   const uint8_t* bitfield, uint32_t max_mip, int snap_to_pixel_centers,
                                   float cone_angle_constant, RefSamplerOut* out);

// The trainable distortion map (Testbed::m_distortion): process-wide state for ref_generate_training_samples / ref_render_nerf (offsets added to the
// ray directions) and ref_compute_cam_gradient (gradient and weight splats, with ref_set_cam_gradient_rng giving the rays' pixels); NULLs = off
void ref_set_distortion(const float* map, int w, int h, float* gradient, float* gradient_weight);
// Error-map importance sampling (testbed.h:733-738): process-wide state for ref_generate_training_samples / ref_compute_loss / ref_compute_cam_gradient.
// Non-NULL CDFs steer the pixel (cdf_x_cond_y + cdf_y) and / or the view (cdf_img) of every ray; a non-NULL error_map receives the rays' mean losses.
void ref_set_error_map(const float* cdf_x_cond_y, const float* cdf_y, const float* cdf_img, int cdf_w, int cdf_h, float* error_map, int map_w, int map_h);
void ref_set_cam_gradient_rng(uint64_t seed, uint64_t skip, const RefImage* images, int snap);
// construct_cdf_2d + construct_cdf_1d + the host-side view CDF (testbed_nerf.cu:2354-2409,3729-3743)
void ref_construct_error_cdfs(uint32_t n_images, uint32_t height, uint32_t width, const float* error_map, float* cdf_x_cond_y, float* cdf_y, float* cdf_img);
// max_level_rand_training (testbed_nerf.cu:1438,1652-1657,1775,1915-1917): a process-wide switch for the two functions around this line; the
// level cut-offs go to max_level [max_samples] / max_level_compacted [max_compacted] when the pointers are not NULL
void ref_set_max_level_rand(int on, float* max_level, float* max_level_compacted);

// --- loss + compositor + gradient (testbed_nerf.cu:186-283,317-387,1340-1355,1660-2012) ---
typedef struct RefLossCfg {
    int loss_type, rgb_activation, density_activation;
    int random_bg_color, linear_colors, snap_to_pixel_centers;
    float background_color[3];
    float loss_scale;          // LOSS_SCALE (128), divided by n_rays inside (testbed_nerf.cu:1903)
    float near_distance;
    float mean_density;        // *mean_density_ptr
    int color_space;           // Testbed::m_color_space (testbed.h:907): 0 Linear (default), 1 SRGB — selects the target formula (testbed_nerf.cu:1801-1818)
    const float* exposure;     // [n_images][3] per-image exposure in stops (cam_exposure_gpu) or NULL = all zero
    float* exposure_gradient;  // [n_images][3] accumulated dL/d exposure (optimize_exposure, :1969-1983) or NULL
    float depth_supervision_lambda;   // nerf.training.depth_supervision_lambda (0 = off; :1848-1856,1934-1939)
    int depth_loss_type;              // nerf.training.depth_loss_type (ELossType, L1 by default: testbed.h:719)
} RefLossCfg;
typedef struct RefLossOut {
    uint16_t* dloss_doutput;     // [max_compacted][4] fp16
    float* coords_compacted;     // [max_compacted][7]
    float* loss;                 // [n_rays]
    uint32_t* numsteps;          // in/out [n_rays][2] -> (compacted_numsteps, compacted_base)
    float* rgb_ray;              // [n_rays][3] (debug: composited prediction)
    uint32_t compacted_counter;  // total compacted steps (may exceed max)
} RefLossOut;
void ref_compute_loss(uint32_t n_rays, const RefAabb* aabb, uint32_t n_rays_total, uint64_t rng_seed, uint64_t rng_skip,
                      uint32_t max_samples_compacted, uint32_t rays_counter, const RefLossCfg* cfg,
                      uint32_t n_images, const RefImage* images, const uint16_t* network_output /*[..][4]*/,
                      const uint32_t* ray_indices, const float* rays, const float* coords, RefLossOut* out);
// compute_cam_gradient_train_nerf (testbed_nerf.cu:2014-2123), extrinsics part: per-image sums (+=) of dL/d(ray origin) and of
// cross(ray direction, dL/d(ray direction)) over the kept rays; numsteps = compacted (count, base), coords / coords_gradient [*][7]
void ref_compute_cam_gradient(uint32_t n_rays, const RefAabb* aabb, uint32_t n_images, uint32_t rays_counter, const uint32_t* ray_indices,
                              const float* rays, const uint32_t* numsteps, const float* coords, const float* coords_gradient,
                              float* cam_pos_gradient, float* cam_rot_gradient);
// fill_rollover[_and_rescale] (tcnn; testbed_nerf.cu:4061-4069) on [n_target][stride]
void ref_fill_rollover_half(uint32_t n_target, uint32_t stride, uint32_t n_input, uint16_t* inout, int rescale);
void ref_fill_rollover_float(uint32_t n_target, uint32_t stride, uint32_t n_input, float* inout);
// NerfCounters control law (testbed_nerf.cu:3583-3609)
uint32_t ref_next_rays_per_batch(uint32_t rays_per_batch, uint32_t target_batch, uint32_t measured_batch);

// --- inference render (testbed_nerf.cu:850-1230,2164-2352,2504-2659) ---
typedef struct RefRenderCfg {
    int32_t res[2];
    float focal_length[2];
    float screen_center[2];
    float camera[12];              // mat4x3
    RefAabb render_aabb, train_aabb;
    float cone_angle_constant;
    float min_transmittance;
    uint32_t max_mip;
    uint32_t spp_index;            // sample_index
    int snap_to_pixel_centers;
    int rgb_activation, density_activation;
    int linear_colors;             // m_nerf.training.linear_colors (shade_kernel_nerf)
    int32_t y0, y1;                // row band [y0,y1) to render (tile sharding)
} RefRenderCfg;
// out rgba [H*W][4] float (linear, premultiplied), depth [H*W]; rows outside [y0,y1) untouched
void ref_render_nerf(const RefRenderCfg* cfg, const RefNerfNetCfg* net, const uint16_t* params, const uint8_t* bitfield,
                     float* rgba, float* depth);

#ifdef __cplusplus
}
#endif
