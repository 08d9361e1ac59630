// Launchers of the NeRF-specific kernels (sampler, occupancy grid, loss/compositor, inference tracer).
// Reference: src/testbed_nerf.cu (kernel-by-kernel citations in nerf_kernels.cu).
#pragma once
#include "nerf_device.cuh"

namespace ngp {

// ---- occupancy grid (testbed_nerf.cu:503-592,635-809,3412-3567) ----
void nerf_mark_untrained_grid(cudaStream_t s, uint32_t n_elements, float* grid, uint32_t n_images, const ImageMeta* metas, bool clear_visible);
void nerf_generate_grid_samples(cudaStream_t s, uint32_t n, Pcg32 rng, uint32_t step, Aabb aabb, const float* grid_in, float* pos_out,
                                uint32_t* indices, uint32_t n_cascades, float thresh);
void nerf_splat_grid_samples(cudaStream_t s, uint32_t n, const uint32_t* indices, const __half* density_raw, float* grid_tmp, int density_activation);
void nerf_ema_grid(cudaStream_t s, uint32_t n_elements, float decay, float* grid, const float* grid_tmp);
// mean over cascade 0 (deterministic two-stage reduction) -> *mean_dev ; bitfield + 7 OR-pooled mips
void nerf_update_bitfield(cudaStream_t s, const float* grid, uint32_t n_cascades, uint8_t* bitfield, float* mean_dev, float* partials /* [1024] */);
// 32^3 coarse cells of the unit cube, bit set if the cell or one of its 26 neighbours holds an occupied fine cell (mip 2 of the bitfield)
void nerf_build_coarse_occupancy(cudaStream_t s, const uint8_t* bitfield, uint32_t* coarse /* [1024] */);

// Device-side step parameters.  The rays of step N+1 need rays_per_batch(N+1), which the control law (NerfCounters::update_after_training,
// testbed_nerf.cu:3583-3609) derives from the sample counters of step N.  Waiting for the HOST to read those counters before the next
// sampler can start costs a round trip on the critical path (measured: the training stream idled 65-75 us per 530 us step,
// profiles/r02_timeline_before.txt).  nerf_control_law evaluates the same arithmetic on the device right behind the loss kernel; the sampler
// for step N+1 is enqueued BEFORE the host knows the result, with grids sized for the bounds, and reads the actual counts from here
// (SamplerBuffers::dyn).  overflow = 1: a count exceeded its bound (or was 0) and nothing was generated — the host, which evaluates the
// same law when the counters arrive, sees that too and samples inline.
struct StepParams { uint32_t n_rays, max_inference, scratch_per_ray, overflow; };
void nerf_control_law(cudaStream_t s, const uint32_t* counters, uint32_t rays_per_batch, uint32_t target_batch_size, uint32_t rays_bound,
                      uint32_t samples_bound, StepParams* out_dev);

// ---- training sampler (testbed_nerf.cu:1382-1658), deterministic ray-order compaction ----
// Error-map importance sampling (testbed.h:733-738; testbed_nerf.cu:1240-1338,2354-2409,3657-3751): piecewise-constant CDFs over an
// error_map-sized lattice per view (columns given the row, rows) and over the views; null pointers = uniform sampling
struct ErrorMapCdf {
    const float* cdf_x_cond_y;   // [n_images][res[1]][res[0]]  (sample_focal_plane_proportional_to_error)
    const float* cdf_y;          // [n_images][res[1]]
    const float* cdf_img;        // [n_images]                  (sample_image_proportional_to_error)
    int res[2];
};
struct SamplerBuffers {
    // per generated ray slot [n_rays]
    uint32_t* numsteps_all;     // pass-1 step count of ray i (0: no sample / masked)
    uint32_t* block_steps;      // [n_rays / 8 + 1] per-block totals (a block holds 256 / lanes_per_ray rays) -> exclusive prefix after the scan
    uint32_t* block_rays;       // [n_blocks]
    float* ray_tmp;             // [n_rays][8]: o(3), d unnormalised(3), startt, cone_angle (pass 1 -> pass 2)
    // compacted outputs
    uint32_t* ray_indices;      // [n_rays]
    float* rays;                // [n_rays][6] unnormalised o,d
    uint32_t* numsteps;         // [n_rays][2] (numsteps, base)
    float* coords;              // [max_samples][7]
    uint32_t* counters;         // [8]: [0] ray_counter, [1] numsteps_counter (total, may exceed max_samples), [2] compacted counter, [3] pool slots taken, [4] candidate rays
    float2* sample_scratch;     // [n_rays][scratch_per_ray] (t, dt) of the samples found by the counting walk
    uint32_t scratch_per_ray;   // window per ray; rays that outgrow it continue in a pool slot (NERF_STEPS entries each)
    float2* sample_pool; uint32_t pool_slots; uint32_t* overflow_slot;   // [pool_slots][NERF_STEPS]; [n_rays] slot of ray i or NO_SLOT; counters[3] = slots taken
    // ray classification (group marchers): rays that cannot touch an occupied cell are never walked
    const uint32_t* coarse;     // 32^3-bit dilated coarse occupancy of the unit cube (nerf_build_coarse_occupancy) or nullptr: every ray is walked
    uint32_t* cand_list;        // [n_rays] indices of the rays to walk; counters[4] = how many
    uint32_t lanes_per_ray;     // 0: choose from n_rays (nerf_sampler_lanes_per_ray); 1 / 4 / 8 / 32: forced (tests compare the variants)
    // non-null (group marchers only): n_rays / max_samples / scratch_per_ray come from device memory, the n_rays and max_samples
    // arguments of nerf_generate_training_samples are then upper bounds that size the grids
    const StepParams* dyn;
    // max_level_rand_training (testbed_nerf.cu:1438): every ray that passes the mask test draws one more random number (its level cut-off)
    // between the pixel position and the motion-blur time, which shifts everything drawn after it
    uint32_t max_level_rand;
    ErrorMapCdf cdf;            // all null: uniform pixels, views in ray order
    DistortionMap distortion;   // data null: none (optimize_distortion off and never trained)
};
uint32_t nerf_sampler_lanes_per_ray(uint32_t n_rays);
// max_level of the un-compacted samples (testbed_nerf.cu:1438,1652-1657): ray i's cut-off = 2 * (the random number after its pixel position), written
// to max_level[base .. base + numsteps) of the samples the sampler kept for it.  Run behind nerf_generate_training_samples with the same rng.
void nerf_fill_max_level(cudaStream_t s, uint32_t n_rays_bound, Pcg32 rng, uint32_t n_images, const ImageMeta* metas, bool snap_to_pixel_centers,
                         uint32_t ray_offset, uint32_t n_rays_global, const SamplerBuffers& sb, float* max_level);
// The group marchers run in two phases that may be launched apart.  WALK (classify, counting walk with its sample log, scan of the block
// totals) needs the ray count, the rng and the bitfield but NOT max_samples; EMIT (ray compaction under the max_samples cap, log -> coordinates)
// is the cheap rest.  The testbed starts the WALK of step N + 2 while step N + 1 is still running (Testbed::launch_walk_ahead).
enum SamplerPhase : int { SAMPLER_ALL = 0, SAMPLER_WALK = 1, SAMPLER_EMIT = 2 };
void nerf_generate_training_samples(cudaStream_t s, uint32_t n_rays, Aabb aabb, uint32_t max_samples, uint32_t n_rays_total, Pcg32 rng,
                                    uint32_t n_images, const ImageMeta* metas, const uint8_t* bitfield, uint32_t max_mip,
                                    bool snap_to_pixel_centers, float cone_angle_constant, uint32_t ray_offset, uint32_t n_rays_global,
                                    const SamplerBuffers& b, int phase = SAMPLER_ALL);

// ---- loss + compositor + gradient (testbed_nerf.cu:1660-2012) ----
struct LossConfig {
    int loss_type, rgb_activation, density_activation;
    int random_bg_color, linear_colors, snap_to_pixel_centers;
    float background_color[3];
    float loss_scale;       // LOSS_SCALE; divided by the (global) ray count inside
    float near_distance;
    int color_space;                // Testbed::m_color_space: 0 Linear (default) / 1 SRGB selects the target formula (testbed_nerf.cu:1801-1818)
    const float* exposure;          // [n_images][3] per-image exposure in stops (cam_exposure_gpu) or nullptr = all zero (scale exactly 1)
    float* exposure_gradient;       // [n_images][3] accumulated dL/d exposure (optimize_exposure, :1969-1983) or nullptr
    float depth_supervision_lambda; // > 0: rays of views with a depth map add lambda * dL_depth/d(depth of the ray) to the density gradient (:1848-1856,1934-1939)
    int depth_loss_type;            // loss of the depth term (ELossType, L1 by default)
    int max_level_rand;             // max_level_rand_training (:1775): one more draw in the replayed rng stream, the value goes to max_level_compacted
};
constexpr uint32_t RAY_SCRATCH = 16;   // + pixel position (2), view pdf, pixel pdf of the ray (error-map sampling)
struct LossBuffers {
    const __half* network_output;   // [*][4]
    __half* dloss_doutput;          // [max_compacted][4]
    float* coords_compacted;        // [max_compacted][7]
    float* loss;                    // [n_rays]
    float* ray_scratch;             // [n_rays][RAY_SCRATCH]: rgb_ray(3), target(3), image index, level cut-off, depth of the ray, depth-loss gradient
    uint32_t* compacted_numsteps;   // [n_rays] pass-1 compacted step counts
    float* t_before;                // [max_samples] transmittance in front of each sample that survives (pass 1 -> pass 2)
    // optional: encoder outputs of the un-compacted samples (blk8, enc_width columns) are carried along with the samples that
    // survive compaction, so the training forward pass does not gather the hash grid a second time for the same points
    const __half* enc_in; __half* enc_out; uint32_t enc_width;
    float* max_level_compacted;     // [max_compacted] level cut-off of each kept sample (max_level_rand_training, :1915-1917) or nullptr
    float* error_map; int error_map_res[2];   // [n_images][res[1]][res[0]] mean ray loss splatted bilinearly at the ray's pixel (:1869-1900) or nullptr
};
// construct_cdf_2d + construct_cdf_1d (testbed_nerf.cu:2354-2409): error map -> per-row column CDFs, per-view row CDFs (both floored by MIN_PDF = 0.01
// of uniform) and the un-normalised per-view mass in cdf_img (the caller turns it into the view CDF on the host, :3729-3743)
void nerf_construct_error_cdfs(cudaStream_t s, uint32_t n_images, uint32_t height, uint32_t width, const float* error_map, float* cdf_x_cond_y, float* cdf_y, float* cdf_img);
// compute_cam_gradient_train_nerf (testbed_nerf.cu:2014-2123), the extrinsics part: per kept ray the input gradients of its compacted samples
// (coords_gradient [*][7]: position cols 0..2, direction cols 4..6) are summed into dL/d(ray origin) and dL/d(ray direction); the origin part is
// added to cam_pos_gradient[image], cross(ray direction, direction part) to cam_rot_gradient[image] (angle-axis, log space).  Run behind the
// backward pass; numsteps are the COMPACTED (count, base) pairs the loss kernel left.
// distortion_gradient / distortion_gradient_weight (optional, [res[1]][res[0]][2] fp32 each, lattice of sb.distortion): the ray's direction gradient,
// projected onto the plane normal to the ray and rotated into the view's image plane, splatted bilinearly at the ray's uv together with the splat
// weights (deposit_image_gradient, common_device.cuh:124-156; testbed_nerf.cu:2088-2100)
void nerf_compute_cam_gradient(cudaStream_t s, uint32_t n_rays_bound, Aabb aabb, Pcg32 rng, bool snap_to_pixel_centers, const ImageMeta* metas, uint32_t n_images,
                               uint32_t ray_offset, uint32_t n_rays_global, const SamplerBuffers& sb, const float* coords_compacted, const float* coords_gradient,
                               float* cam_pos_gradient, float* cam_rot_gradient, float* distortion_gradient = nullptr, float* distortion_gradient_weight = nullptr);
// the camera update of the distortion map (testbed_nerf.cu:3811-3818): gradient / weight where the weight is positive (safe_divide, :2411-2417), then
// one step of tiny-cuda-nn's Adam on fp32 parameters that are all "non-matrix" (entries with a zero gradient are skipped, no L2 term; recalled)
void nerf_distortion_step(cudaStream_t s, uint32_t n_params, float* params, float* gradient, const float* gradient_weight, float* first_moment, float* second_moment,
                          uint32_t* steps, float learning_rate, float beta1, float beta2, float epsilon, float loss_scale);
void nerf_compute_loss(cudaStream_t s, uint32_t n_rays, Aabb aabb, uint32_t n_rays_total, Pcg32 rng, uint32_t max_samples_compacted,
                       const LossConfig& cfg, uint32_t n_images, const ImageMeta* metas, const float* mean_density_dev,
                       uint32_t ray_offset, uint32_t n_rays_global, const SamplerBuffers& sb, const LossBuffers& lb);
// optimize_exposure (testbed_nerf.cu:3831-3858), on the device: one AdamOptimizer<vec3> step per image (adam_optimizer.h:143-150: lr sqrt(1 - b2^t) / (1 - b1^t),
// beta1 0.9, beta2 0.99, epsilon 1e-8) on gradient * per_camera_loss_scale + l2_reg * exposure, then the mean over the images is subtracted
void nerf_exposure_step(cudaStream_t s, uint32_t n_images, const float* gradient, float* exposure, float* first_moment, float* second_moment, uint32_t iter,
                        float learning_rate, float l2_reg, float per_camera_loss_scale);
// fill_rollover_and_rescale + fill_rollover (tcnn; testbed_nerf.cu:4061-4069) on dL [n_target][4] fp16 and coords [n_target][7]
void nerf_fill_rollover(cudaStream_t s, uint32_t n_target, const uint32_t* n_input_dev, __half* dloss, float* coords, __half* enc = nullptr, uint32_t enc_width = 0, float* max_level_compacted = nullptr);
// sum of loss[0..n) -> *out_dev (deterministic)
void reduce_sum_f32(cudaStream_t s, const float* in, uint32_t n, float* out_dev, float* partials /* [1024] */);

// ---- inference tracer (testbed_nerf.cu:850-1230,2164-2352,2417-2693) ----
// ERenderMode (common.h:110-121).  Normals needs the network's input gradient and is rejected by the host; Distortion and Slice have no
// NeRF-specific compositing in the reference either and shade like Shade (Slice: no depth write).
enum RenderMode : int { RM_AO = 0, RM_SHADE = 1, RM_NORMALS = 2, RM_POSITIONS = 3, RM_DEPTH = 4, RM_DISTORTION = 5, RM_COST = 6, RM_SLICE = 7, RM_ENCODING_VIS = 9 };
struct NerfPayload {   // nerf.h:32-40
    float origin[3]; float dir[3]; float t; float max_weight; uint32_t idx; uint16_t n_steps; uint8_t alive; uint8_t pad;
};
struct RaysSoa { float* rgba; float* depth; NerfPayload* payload; };   // rgba [n][4]
struct RenderCamera {
    int res[2]; float focal_length[2]; float screen_center[2]; float camera[12];
    int y0, y1;     // rows y0, y0 + y_stride, ... < y1 are rendered by this call (tile sharding: a band, or every world-th row); local ray index = x + W*((y - y0) / y_stride)
    int y_stride;
    int lens_mode; float lens_params[4];   // m_nerf.render_lens (render_with_lens_distortion): Perspective or OpenCV
    DistortionMap distortion;              // the trained distortion map, with render_with_lens_distortion (testbed_nerf.cu:2776-2778); data null: none
};
// Loop control of the wavefront tracer lives on the device (the reference reads n_alive back and synchronises every iteration,
// testbed_nerf.cu:2540-2560): iteration `it` reads ctl[(it - 1) & 1] and its compaction writes ctl[it & 1] plus a copy into host-mapped
// memory, which the host uses only to shrink the launch bounds and to notice the end — a few iterations late, never waiting for the GPU.
struct TraceCtl {
    uint32_t n_alive;        // rays still marching after this iteration's compaction
    uint32_t n_steps;        // samples per ray this iteration: clamp(TRACE_TARGET_SAMPLES / n_alive, 1, cap) (:2574-2578), cap raised when few rays are left
    uint32_t n_elements;     // n_alive * n_steps rounded up to the batch granularity: rows of the network call
    uint32_t current_step;   // first step index of this iteration (the `i` of the reference's loop; Cost mode reports it)
};
constexpr uint32_t TRACE_MAX_STEPS = 8, TRACE_MIN_STEPS = 1;   // :97-100
// samples per iteration the schedule aims for.  The reference uses 2 Mi (:2574), sized for the memory of a desktop GPU; a 3840 x 2160 frame then
// spends its first 20 iterations at one sample per ray, paying the compaction, seven launches and every per-ray load once per sample.  8 Mi
// (0.5 GB of sample rows: nothing on a 180 GB part) starts such a frame at 5 samples per ray: 40 -> 19 iterations, 17.9 -> 15.0 ms
// (16 Mi: 15.2 ms — the samples wasted behind a ray's last one grow with the step count).  NGP_TRACE_TARGET_SAMPLES overrides (experiments).
constexpr uint32_t TRACE_TARGET_SAMPLES = 8u * 1024u * 1024u;
constexpr uint32_t TRACE_MAX_STEPS_TAIL = 32;     // see trace_steps()
// samples per ray and iteration.  The reference's schedule up to 8; beyond that only when so few rays are alive that an iteration is bound
// by launch latency, not throughput (<= 131072: 16, <= 32768: 32) — a ray's result does not depend on the schedule (the compositor consumes
// samples in order and stops at the same one), only the number of iterations does; `capacity` = rows the sample buffers hold
NGP_HD uint32_t trace_steps(uint32_t n_alive, uint32_t capacity, uint32_t target_samples) {
    if (n_alive == 0) return 0;
    const uint32_t cap = n_alive <= 32768u ? TRACE_MAX_STEPS_TAIL : (n_alive <= 131072u ? 16u : TRACE_MAX_STEPS);
    uint32_t k = target_samples / n_alive;
    k = k < TRACE_MIN_STEPS ? TRACE_MIN_STEPS : (k > cap ? cap : k);
    const uint32_t fit = capacity / n_alive;
    return k > fit ? (fit > 0 ? fit : 1u) : k;
}
// coarse (optional): 32^3 dilated occupancy bits of the unit cube (nerf_build_coarse_occupancy); rays that cannot touch an occupied cell
// are not walked and look-ups inside empty coarse cells are not made — same results, see the kernels
void nerf_init_rays(cudaStream_t s, const RenderCamera& cam, uint32_t sample_index, bool snap_to_pixel_centers, Aabb render_aabb,
                    const uint8_t* bitfield, uint32_t max_mip, float cone_angle, RaysSoa rays, float* frame_rgba, float* frame_depth, float near_distance = 0.f,
                    const uint32_t* coarse = nullptr);
// alive -> dst (compacted, order preserving), finished with alpha > 0.001 -> hit ; counters[0] = n_alive, counters[1] += n_hit.
// n_bound >= prev->n_alive sizes the launch; writes *cur (and *host_copy, host-mapped) for the iteration that follows
void nerf_compact_rays(cudaStream_t s, uint32_t n_bound, RaysSoa src, RaysSoa dst, RaysSoa hit, uint32_t* counters, uint32_t* scan_scratch,
                       const TraceCtl* prev, TraceCtl* cur, TraceCtl* host_copy, uint32_t capacity, uint32_t target_samples);
void nerf_generate_next_inputs(cudaStream_t s, uint32_t n_bound, const TraceCtl* ctl, Aabb render_aabb, Aabb train_aabb, RaysSoa rays,
                               float* network_input /*[n_alive*k][7]*/, const uint8_t* bitfield, uint32_t max_mip, float cone_angle, const uint32_t* coarse = nullptr);
void nerf_composite(cudaStream_t s, uint32_t n_bound, const TraceCtl* ctl, Aabb train_aabb, const float* camera, RaysSoa rays,
                    const float* network_input, const __half* network_output, int rgb_activation, int density_activation,
                    float min_transmittance, int render_mode = RM_SHADE, float depth_scale = 1.f);
// n_hit_dev: number of finished rays (counters[1]); n_bound sizes the launch
void nerf_shade(cudaStream_t s, uint32_t n_bound, const uint32_t* n_hit_dev, RaysSoa hit, bool linear_colors, const RenderCamera& cam, float* frame_rgba, float* frame_depth,
                int render_mode = RM_SHADE);

}  // namespace ngp
