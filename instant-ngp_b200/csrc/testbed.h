// Host driver of the hot path: the slice of ngp::Testbed that owns training and inference
// (reference: include/neural-graphics-primitives/testbed.h:64-516, src/testbed.cu:3903-4151,4285-4370,
// src/testbed_nerf.cu:3412-3609,3611-3862,3867-4152, src/testbed_image.cu:214-285,465-527).
#pragma once
#include "network.h"
#include "nerf_kernels.h"
#include "image_kernels.h"
#include "sdf_kernels.h"
#include "frame_output.h"
#include "comm.h"
#include <memory>
#include <array>
#include <vector>

namespace ngp {

enum TestbedMode : int { MODE_TB_NERF = 0, MODE_TB_SDF = 1, MODE_TB_IMAGE = 2 };

struct HostImage {   // one training view handed over by the caller (host memory)
    const void* pixels; int data_type; int res[2]; float focal_length[2]; float principal_point[2]; int lens_mode; float lens_params[4]; float xform[12];
};

struct TrainStats {
    uint32_t training_step; float loss; uint32_t rays_per_batch, measured_batch_size, measured_batch_size_before_compaction, n_rays_total;
    float learning_rate; float density_grid_mean;
};

bool host_has_mask_pixels(const void* rgba8_pixels, size_t n_pixels);   // MaskScanPool (testbed.cu)
class Testbed {
public:
    explicit Testbed(int mode);
    ~Testbed();

    // reload_network_from_json (testbed.cu:281-292) -> reset_network (testbed.cu:3903-4151)
    void reload_network_from_json(const std::string& json, const std::string& config_base_path = std::string());
    // reload_network_from_file / find_network_config (testbed.cu:228-244,275-309): <path>, else <root_dir>/configs/<mode>/<path>; "parent" configs merged
    void reload_network_from_file(const std::string& path = std::string());
    std::string find_network_config(const std::string& path) const;
    static const char* builtin_base_config(int mode);
    // Testbed::m_max_level_rand_training (testbed.h:573, python_api.cu:498): every training ray draws a level cut-off 2u, u uniform in [0,1): its samples
    // are encoded with the hash-grid levels below cut-off * n_levels only (half of the rays keep every level)
    bool max_level_rand_training = false;
    // nerf.training.sample_focal_plane_proportional_to_error / sample_image_proportional_to_error (testbed.h:733-738; python_api.cu:624-625): pixels
    // (half of them) and / or views are drawn from CDFs built from an error map, the mean loss of past rays splatted per view; the map is
    // accumulated while either switch is on and turned into CDFs every n_steps_between_error_map_updates steps (128, growing by 1.5x each time)
    bool sample_focal_plane_proportional_to_error = false, sample_image_proportional_to_error = false;
    uint32_t n_steps_between_error_map_updates = 128;
    // nerf.training.depth_supervision_lambda / depth_loss_type (testbed.h:726,699; python_api.cu:616,635): weight and loss of the depth term that views
    // with a depth map add to the density gradient (testbed_nerf.cu:1848-1856,1934-1939)
    float depth_supervision_lambda = 0.f; int depth_loss_type = LOSS_L1;
    // nerf.training.optimize_extrinsics / extrinsic_l2_reg / extrinsic_learning_rate (testbed.h:704,708,709; python_api.cu:624,631,632): per-view
    // position and rotation offsets trained from the network's input gradients (testbed_nerf.cu:2014-2123,3780-3808,2984-3023)
    struct VecAdam { uint32_t iter = 0; float m1[3] = {0, 0, 0}, m2[3] = {0, 0, 0}, var[3] = {0, 0, 0}; };   // state of one vec3 optimiser
    bool optimize_extrinsics = false; float extrinsic_l2_reg = 1e-4f, extrinsic_learning_rate = 1e-3f;
    // nerf.training.optimize_distortion (testbed.h:723; python_api.cu:623): a lattice of 2-D image-plane offsets shared by all views (config
    // "distortion_map": resolution 32 x 32 and its own optimizer, testbed.cu:4066-4077), added to every ray direction and trained from the ray-direction
    // gradients at the camera-update cadence (testbed_nerf.cu:2088-2100,3811-3818)
    bool optimize_distortion = false;
    // nerf.training.get_camera_extrinsics(frame_idx, convert_to_ngp) (python_api.cu:594-603; testbed_nerf.cu:2925-2932): the view's pose as trained
    void get_camera_extrinsics(int frame_idx, bool convert_to_ngp, float* out12) const;
    // NerfDataset::set_training_image's depth half (nerf_loader.cu:943-959): host depth map of view frame_idx (depth_type 0: uint16, 1: float32), stored
    // as value * depth_scale; depth_scale <= 0 or a null pointer: a map of zeros ("no depth for this image"); depth_scale < 0 AND null: the map is dropped
    void set_training_depth(cudaStream_t s, int frame_idx, const void* depth_host, int depth_type, float depth_scale);
    std::string root_dir;                                   // Testbed::m_root_dir (testbed.h:978): where configs/<mode>/ is looked for
    std::string network_config_path = "base.json";          // Testbed::m_network_config_path (testbed.h:1004)
    // load_training_data equivalents with host buffers (NerfDataset: nerf_loader.h:110-171; load_nerf_post testbed_nerf.cu:3046-3112)
    void set_nerf_dataset(cudaStream_t s, uint32_t n_images, const HostImage* images, int aabb_scale);
    void set_image(cudaStream_t s, const float* rgba_host, int w, int h);
    void set_image_rgba8(cudaStream_t s, const uint8_t* rgba8_host, int w, int h);
    // override_sdf_training_data (python_api.cu:69-99): n host (point, signed distance) pairs in mesh space, moved into the unit
    // cube with the mesh's raw bounding box and scale; from then on train() runs train_sdf on them (testbed_sdf.cu:1289-1311)
    void override_sdf_training_data(cudaStream_t s, const float* points, const float* distances, uint32_t n, const float* raw_aabb_min,
                                    const float* raw_aabb_max, float mesh_scale);
    // nerf.training.set_camera_intrinsics / set_camera_extrinsics (testbed_nerf.cu:2873-2919): one view's metadata, host + device copy.
    // cx, cy: pixels when >= 0, already-normalised principal point when negative; k1 k2 p1 p2 all zero = Perspective.
    // convert_to_ngp: `camera_to_world` is a NeRF-convention mat4x3 (column major) and goes through nerf_matrix_to_ngp with the
    // scale / offset the data set was loaded with (0.33 / 0.5 for views handed over in memory)
    void set_camera_intrinsics(cudaStream_t s, int frame_idx, float fx, float fy, float cx, float cy, float k1, float k2, float p1, float p2);
    void set_camera_extrinsics(cudaStream_t s, int frame_idx, const float* camera_to_world, bool convert_to_ngp);
    // nerf.training.set_camera_extrinsics_rolling_shutter (python_api.cu:606-617, testbed_nerf.cu:2896-2915): the pose at the start and at the end of the
    // exposure + (A, B, C, D): a ray's camera is interpolated at pixel time A + B u + C v + D motionblur_time (common_device.cuh:401-408)
    void set_camera_extrinsics_rolling_shutter(cudaStream_t s, int frame_idx, const float* c2w_start, const float* c2w_end, const float* rolling_shutter4, bool convert_to_ngp);
    // NerfDataset::scale / offset / aabb_scale / xforms[i] (nerf_loader.h:73-171) of the loaded data set: the poses as given, before trained offsets
    const int* error_map_res() const { return m_error_map_res; }
    const int* error_cdf_res() const { static const int none[2] = {0, 0}; return m_cdf_valid ? m_cdf_res : none; }
    float dataset_scale() const { return m_ds_scale; }
    const float* dataset_offset() const { return m_ds_offset; }
    int dataset_aabb_scale() const { return m_aabb_scale; }
    const float* given_xform(uint32_t i) const { return i < m_base_xforms.size() ? m_base_xforms[i].data() : nullptr; }   // [24]: start, end
    void set_dataset_scale_offset(float scale, const float* offset3) { m_ds_scale = scale; for (int k = 0; k < 3; ++k) m_ds_offset[k] = offset3[k]; }
    // Testbed::reset_network(clear_density_grid) (python_api.cu "reset"): fresh parameters, optimizer state, counters
    void reset(bool reset_density_grid);
    // nerf.training.set_image (python_api.cu:49-67): replace the pixels of one training view (same resolution / type),
    // host -> device; asynchronous with respect to the host when the source is pinned.  RGBA8 views are copied on an internal
    // copy stream, ordered after everything already queued on `s` and before the next training step's loss kernel, so the
    // upload runs beside the start of that step; the host buffer must stay valid until the next train call has returned
    void set_training_image(cudaStream_t s, uint32_t frame_idx, const void* pixels_host);

    // Testbed::train (testbed.cu:4285-4370): one step.  train_begin computes gradients, train_end runs the optimizer and
    // the NerfCounters control law; train = begin + end.  (DP: all-reduce gradient_buffer() between the two.)
    void train(cudaStream_t s, uint32_t batch_size);
    void train_begin(cudaStream_t s, uint32_t batch_size);
    void train_end(cudaStream_t s, uint32_t measured_override, uint32_t measured_before_override);
    // The host half of a NeRF step's end (NerfCounters::update_after_training: wait for the two counters, control law, next prefetch).  When the
    // next step's sampler is already in flight behind the device-side control law, nothing on the device waits for it, so train_end leaves it
    // pending and the host returns to the caller at once; it runs at the start of the next step or when anything else looks at the testbed
    // (every C-ABI entry point except set_training_image calls finish_pending_step first).  deferred_step_finish = false: always in train_end.
    void finish_pending_step();
    bool deferred_step_finish = true;
    uint32_t training_step() const { return m_training_step; }                     // host counter: valid while a step's host half is pending
    uint64_t measured_total() const { return m_measured_total; }                  // sums of the per-step counts since the testbed was made
    uint64_t measured_before_total() const { return m_measured_before_total; }
    void set_shard(uint32_t rank, uint32_t world) { m_rank = rank; m_world = world; }
    // local sample counts of the step begun by train_begin, available as soon as its loss kernel has run (the backward pass is
    // still in flight): lets a data-parallel host exchange them while the device works
    void pending_counters(uint32_t* measured, uint32_t* measured_before);
    // attach an NCCL communicator (collective: every rank calls it with the same id): from then on train() is the data-parallel
    // step — rays sharded by rank, sample counters and gradients all-reduced from inside the step, the gradient slices one hash
    // level chunk at a time on a communication stream beside the remaining backward kernels
    void comm_init(const void* id128, uint32_t rank, uint32_t world, const char* nccl_lib_path);

    // render_nerf (testbed_nerf.cu:2752-2871) into a host or device float4 buffer, rows [y0,y1), averaged over spp
    void render(cudaStream_t s, const float* camera_matrix, const float* focal_length, const float* screen_center, int w, int h, int spp, int y0, int y1,
                float* rgba_out, bool out_is_device, float* depth_out);
    // Testbed::render_to_cpu -> render_frame (python_api.cu:127-205, testbed.cu:4480-4760), headless: rows [y0,y1) of a w x h frame seen by the
    // testbed's OWN camera (camera / relative_focal_length / fov_axis / zoom / screen_center below), spp samples accumulated in
    // `color_space`, then background + exposure + tonemap curve + output colour space (sRGB when to_srgb), then the ground-truth overlay
    // when render_ground_truth is set (with ground_truth_alpha >= 1 the network is not traced at all, as in the reference)
    void render_frame(cudaStream_t s, int w, int h, int spp, bool to_srgb, int y0, int y1, float* rgba_out, bool out_is_device);
    // view state helpers (testbed.cu:789-880,3810-3824)
    void reset_camera();
    void set_fov(float degrees);
    float fov() const;
    void set_nerf_camera_matrix(const float* nerf_mat4x3_col_major);
    void set_camera_to_training_view(int trainview);
    void apply_dataset_view_hints(bool has_render_aabb, const float* render_aabb6, const float* up3);   // load_nerf_post :3096-3111
    float compute_image_mse(cudaStream_t s, bool quantize);   // testbed_image.cu:465-527
    // save_snapshot / load_snapshot (testbed.cu:4873-5057): the network config plus a "snapshot" object (fp16 inference parameters,
    // fp16 density grid, counters, optionally the optimizer state) as MessagePack; ".ingp" = the same inside a zlib stream
    void save_snapshot(cudaStream_t s, const std::string& path, bool include_optimizer_state, bool compress);
    void load_snapshot(cudaStream_t s, const std::string& path);
    // the "mode" a snapshot file was saved in (MODE_TB_*; testbed.cu:4950-4955: files without the key are NeRF snapshots), read without a testbed
    static int snapshot_mode(const std::string& path);

    TrainStats stats() const;
    Trainer* trainer() { return m_trainer.get(); }
    NerfNetwork* nerf_network() { return m_nerf_net.get(); }
    uint64_t n_params() const { return m_trainer ? m_trainer->n_params : 0; }
    const uint8_t* density_grid_bitfield() const { return m_bitfield.as<uint8_t>(); }
    const float* density_grid() const { return m_density_grid.as<float>(); }
    uint32_t max_cascade() const { return m_max_cascade; }
    // parity hook: internal per-step buffers by name (tests compare them with the oracle)
    void debug_buffer(const std::string& name, void** ptr, uint64_t* bytes);

    // runtime knobs (pyngp def_readwrite members: python_api.cu:452-537,611-695)
    bool random_bg_color = true, linear_colors = false, shall_train_encoding = true, shall_train_network = true;
    // two different members of the reference: nerf.training.snap_to_pixel_centers (training rays, default true, testbed.h:773) and
    // Testbed::m_snap_to_pixel_centers (render rays, default false, python_api.cu:478)
    bool train_snap_to_pixel_centers = true, snap_to_pixel_centers = false;
    float near_distance = 0.1f, density_grid_decay = 0.95f, render_min_transmittance = 0.01f, cone_angle_constant = 0.0f;
    float background_color[4] = {0.f, 0.f, 0.f, 1.f};   // sRGB RGBA (testbed.h:936); training composites on .rgb, the output stage uses all four
    // view + output-stage state (testbed.h:560-610,904-936)
    float camera[12] = {1, 0, 0, 0, -1, 0, 0, 0, -1, 0.5f, 0.5f, 2.0f};   // m_camera, column-major mat4x3 (reset_camera)
    float relative_focal_length[2] = {1.f, 1.f}, screen_center[2] = {0.5f, 0.5f}, zoom = 1.f, scale = 1.5f, up_dir[3] = {0.f, 1.f, 0.f};
    int fov_axis = 1, training_view = 0;
    float exposure = 0.f, ground_truth_alpha = 1.f, render_near_distance = 0.f;
    int color_space = CS_LINEAR, tonemap_curve = TM_IDENTITY, render_mode = RM_SHADE;
    // tile-sharded inference: render() / render_frame() produce rows y0, y0 + render_row_stride, ... < y1 (default 1: the contiguous band [y0, y1));
    // with stride = world and y0 = rank every rank gets the same share of every part of the picture, whatever the scene covers
    int render_row_stride = 1;
    bool render_ground_truth = false, render_with_lens_distortion = false;
    int loss_type = LOSS_L2, rgb_activation = ACT_LOGISTIC, density_activation = ACT_EXPONENTIAL;
    bool image_snap_to_pixel_centers = true, image_linear_colors = false;
    int image_random_mode = IMG_RANDOM_STRATIFIED;   // image.random_mode (testbed.h:881)
    uint64_t seed = 1337;
    uint32_t sampler_lanes_per_ray = 0;   // 0 = automatic; 1 / 2 / 4 / 8 / 32 force a marcher variant (parity tests)
    // generate the NEXT step's rays on a second stream while this step's backward pass and optimizer run (the walk depends on
    // the occupancy bitfield and the rng only, not on the weights); same samples either way, see launch_ray_prefetch()
    bool prefetch_rays = true;
    // nerf.training.optimize_exposure / exposure_l2_reg / n_steps_between_cam_updates (testbed.h:712,727,730): per-image exposure (stops) trained
    // with its own Adam every n_steps_between_cam_updates steps from the gradient the loss kernel accumulates; everything stays on the device
    bool optimize_exposure = false; float exposure_l2_reg = 0.0f; uint32_t n_steps_between_cam_updates = 16;
    // general marcher (cascades / cone angle: the walk takes longer than the rest of the step): start the counting walk of step N + 2 during
    // step N + 1, with the ray count PREDICTED to stay what step N + 1 uses (rays_per_batch is a multiple of 256 and moves rarely at steady
    // state); when the control law then says otherwise the walk is discarded and redone.  Same samples either way.
    // OFF by default — measured on fox (profiles/r02_fox_walk_ahead_sweep.json, 5000 steps, ms/step off -> on): 8 lanes per ray 1.38 -> 1.65,
    // 4 lanes 1.62 -> 1.86, 2 lanes 1.99 -> 2.42, 32 lanes 2.01 -> 2.28; 70 % of the predictions hold.  The general marcher is not idle while it
    // waits on its chain — eight lanes each run the whole chain of a round — so a walk that runs beside every kernel of the step slows those
    // kernels by as much as it hides (fused MLP forward 0.03 -> 0.5 ms under it).  Kept as an option with its bit-exactness test.
    bool walk_ahead = false;
    uint32_t walk_ahead_hits = 0, walk_ahead_misses = 0;   // read-only
    // evaluate the rays_per_batch control law on the device as well and enqueue the next step's sampler before the host has read this
    // step's counters (nerf_kernels.h StepParams); off: the sampler of step N+1 is launched by the host after it has read them.  Same results.
    bool device_control_law = true;
    uint32_t prefetch_hits = 0;   // read-only: steps whose rays came from a prefetch
    // data parallel only: reduce-scatter the gradients, run Adam(+EMA) on this rank's 1/world of the entries, all-gather the fp16
    // parameters (instead of all-reduce + the full optimizer on every rank).  Same arithmetic per entry; the EMA (inference)
    // parameters of the other ranks' slices are fetched by sync_inference_params() — a COLLECTIVE every rank must call before
    // any of them renders or saves a snapshot (pyngp does it inside render / save_snapshot).
    bool dp_sharded_optimizer = false;
    void sync_inference_params(cudaStream_t s);
    Aabb aabb{{0, 0, 0}, {1, 1, 1}}, render_aabb{{0, 0, 0}, {1, 1, 1}};
    int render_lens_mode = LENS_PERSPECTIVE; float render_lens_params[4] = {0, 0, 0, 0};   // m_nerf.render_lens (set_camera_to_training_view)
    const std::vector<ImageMeta>& training_views() const { return m_host_metas; }

private:
    void reset_network(bool clear_density_grid = true);
    void upload_view_metadata(cudaStream_t s, uint32_t frame_idx);
    void load_nerf_post(cudaStream_t s, int aabb_scale);                                           // testbed_nerf.cu:3046-3112
    void training_prep_nerf(cudaStream_t s);                                                       // testbed_nerf.cu:4137-4152
    void update_density_grid_nerf(cudaStream_t s, float decay, uint32_t n_uniform, uint32_t n_nonuniform);   // :3412-3536
    void train_nerf_step(cudaStream_t s, uint32_t target_batch_size);                              // :3867-4132
    void train_image_step(cudaStream_t s, uint32_t batch_size, bool get_loss);                     // testbed_image.cu:214-285
    void train_sdf_step(cudaStream_t s, uint32_t batch_size, bool get_loss);                       // testbed_sdf.cu:1289-1311
    struct SamplerKey {   // everything generate_training_samples depends on besides the bitfield and the view poses
        uint32_t step, n_rays, max_inference, target, rank, world, lanes, dataset_version; uint64_t rng_state, rng_inc; float cone; bool snap, max_level_rand; uint32_t cdf_version, distortion_version;
        bool operator==(const SamplerKey& o) const {
            return step == o.step && n_rays == o.n_rays && max_inference == o.max_inference && target == o.target && rank == o.rank && world == o.world &&
                   lanes == o.lanes && dataset_version == o.dataset_version && rng_state == o.rng_state && rng_inc == o.rng_inc && cone == o.cone && snap == o.snap && max_level_rand == o.max_level_rand && cdf_version == o.cdf_version && distortion_version == o.distortion_version;
        }
    };
    SamplerKey sampler_key(uint32_t n_rays, uint32_t max_inference, uint32_t target) const;
    SamplerBuffers sampler_buffers(cudaStream_t s, uint32_t set, uint32_t target, uint32_t n_rays);
    uint32_t max_inference_samples(uint32_t target);
    void launch_sampler(cudaStream_t s, const SamplerKey& key, const SamplerBuffers& sb);
    void launch_ray_prefetch(cudaStream_t s, uint32_t target);

    int m_mode;
    Json m_config; bool m_have_config = false;
    std::unique_ptr<NerfNetwork> m_nerf_net;
    std::unique_ptr<EncodedMlp> m_image_net;
    std::unique_ptr<Trainer> m_trainer;
    uint32_t m_training_step = 0; float m_loss_scalar = 0.f; bool m_get_loss = false; uint32_t m_pending_batch = 0;
    Pcg32 m_rng, m_density_grid_rng;
    uint32_t m_rank = 0, m_world = 1;

    // NeRF dataset + occupancy grid
    float m_density_grid_mean = 0.f;      // host copy, refreshed with the loss every 16th step
    void finish_step(cudaStream_t s, uint32_t target, uint32_t measured_override, uint32_t measured_before_override);
    bool m_finish_pending = false; uint32_t m_finish_target = 0; cudaStream_t m_finish_stream = nullptr;
    uint64_t m_measured_total = 0, m_measured_before_total = 0;
    std::vector<bool> m_frame_has_mask;   // RGBA8 view contains the mask colour 0x00FF00FF (see set_training_image)
    uint32_t m_n_images = 0; int m_aabb_scale = 1; uint32_t m_max_cascade = 0;
    std::vector<float*> m_depth_buffers;                    // per view: device depth map (NerfDataset::depthmemory) or nullptr
    std::vector<void*> m_image_buffers; std::vector<size_t> m_image_bytes; DeviceBuffer m_metas; std::vector<ImageMeta> m_host_metas;
    DeviceBuffer m_density_grid, m_density_grid_tmp, m_bitfield, m_mean, m_partials, m_grid_pos, m_grid_idx, m_grid_out;
    uint32_t m_density_grid_ema_step = 0; bool m_grid_allocated = false;
    DeviceBuffer m_error_map, m_cdf_x_cond_y, m_cdf_y, m_cdf_img; int m_error_map_res[2] = {0, 0}, m_cdf_res[2] = {0, 0};
    uint32_t m_n_steps_since_error_map_update = 0, m_cdf_version = 0; bool m_cdf_valid = false, m_error_map_live = false;
    void update_error_map_cdfs(cudaStream_t s);
    DeviceBuffer m_max_level, m_max_level_compacted;        // per-sample level cut-offs (max_level_rand_training)
    size_t m_grid_elements = 0;                             // cells of the occupancy grid as allocated (128^3 x cascades in use)
    float m_dataset_up[3] = {0.f, 1.f, 0.f};                // NerfDataset::up (nerf_loader.h:88) of the loaded data set
    float m_ds_scale = 0.33f, m_ds_offset[3] = {0.5f, 0.5f, 0.5f};   // NerfDataset::scale / offset (nerf_loader.h:29, nerf_loader.cu:392-393)
    // NerfCounters (testbed.h:433-447)
    uint32_t m_rays_per_batch = 1u << 12, m_n_rays_total = 0, m_measured = 0, m_measured_before = 0;
    // per-step scratch
    // sampler outputs and scratch exist twice: step N reads set N & 1 while the prefetch of step N + 1 fills the other one
    static constexpr uint32_t N_SETS = 3;   // step N reads set N % 3; N + 1's rays are prefetched into the next and the walk of N + 2 may already fill the third
    DeviceBuffer m_numsteps_all[N_SETS], m_block_steps[N_SETS], m_block_rays[N_SETS], m_ray_tmp[N_SETS], m_ray_indices[N_SETS], m_rays[N_SETS], m_numsteps[N_SETS],
        m_coords[N_SETS], m_counters[N_SETS], m_sample_scratch[N_SETS], m_sample_pool[N_SETS], m_overflow_slot[N_SETS], m_cand_list[N_SETS];
    DeviceBuffer m_coarse;   // dilated 16^3 coarse occupancy (ray classification), rebuilt with the bitfield
    DeviceBuffer m_mlp_out, m_dloss, m_coords_compacted, m_loss, m_ray_scratch, m_compacted_numsteps, m_loss_sum, m_t_before;
    cudaStream_t m_aux_stream = nullptr; cudaEvent_t m_ev_counters = nullptr, m_ev_prefetch = nullptr, m_ev_after_mlp = nullptr;
    uint32_t* m_host_counters = nullptr;   // pinned: [0..3] counters, [4] loss sum, [5..6] all-reduced (before, measured)
    bool m_inference_stale = false;       // sharded optimizer: inference_params outside this rank's slice are behind
    bool sharded_step() const;            // dp_sharded_optimizer && a communicator && the vector splits evenly over the ranks
    std::unique_ptr<Comm> m_comm; cudaStream_t m_comm_stream = nullptr; cudaEvent_t m_ev_slice = nullptr, m_ev_comm_done = nullptr, m_ev_dp_counters = nullptr;
    DeviceBuffer m_dp_counters;
    cudaStream_t m_copy_stream = nullptr; cudaEvent_t m_ev_copy_gate = nullptr, m_ev_copy_done = nullptr; bool m_copy_pending = false;
    bool m_prefetch_inflight = false;     // a prefetch (adopted or not) may still be writing its buffer set: wait for m_ev_prefetch before reuse
    bool m_early_prefetch = false; uint32_t m_early_rays_bound = 0, m_early_samples_bound = 0, m_early_dataset_version = 0;   // sampler of the next step enqueued from train_nerf_step
    DeviceBuffer m_step_params[N_SETS]; cudaEvent_t m_ev_after_loss = nullptr;
    // walk-ahead (general marcher only, see launch_walk_ahead): the counting walk of step N + 2, started with a PREDICTED ray count
    bool m_walk_ahead_valid = false, m_walk_ahead_inflight = false; SamplerKey m_walk_ahead_key{}; Pcg32 m_walk_ahead_rng{0}; cudaEvent_t m_ev_walk_ahead = nullptr;
    void launch_walk_ahead(uint32_t target);
    bool adopt_walk_ahead(uint32_t target);
    void launch_early_prefetch(cudaStream_t s, uint32_t target, uint32_t rays_bound, uint32_t samples_bound);
    bool m_prefetch_valid = false, m_all_images_u8 = false; SamplerKey m_prefetch_key{}; uint32_t m_dataset_version = 0, m_last_set = 0;
    // camera extrinsics optimisation: the poses as given (start, end), the host-side optimiser state per view (AdamOptimizer<vec3> for the position
    // offset, RotationAdamOptimizer for the rotation offset: adam_optimizer.h:143-301), the device gradient accumulators and dL/d(coords)
    std::vector<std::array<float, 24>> m_base_xforms; std::vector<VecAdam> m_cam_pos_offset, m_cam_rot_offset;
    DeviceBuffer m_cam_pos_grad, m_cam_rot_grad, m_coords_gradient;
    DeviceBuffer m_distortion_map, m_distortion_grad, m_distortion_weight, m_distortion_m1, m_distortion_m2, m_distortion_steps;
    int m_distortion_res[2] = {32, 32}; std::unique_ptr<Optimizer> m_distortion_opt; bool m_distortion_live = false; uint32_t m_distortion_version = 0;
    DistortionMap distortion_view() const { DistortionMap d{}; if (m_distortion_live) { d.data = (const float*)m_distortion_map.ptr; d.res[0] = m_distortion_res[0]; d.res[1] = m_distortion_res[1]; } return d; }
    void ensure_distortion_map(cudaStream_t s);
    void distortion_step(cudaStream_t s);
    void effective_xform(uint32_t i, float* start12, float* end12);
    void update_transforms(cudaStream_t s, uint32_t first, uint32_t last);   // NerfDataset xforms + offsets -> the poses the kernels see (testbed_nerf.cu:2984-3023)
    void extrinsics_step(cudaStream_t s);
    DeviceBuffer m_cam_exposure, m_cam_exposure_grad, m_cam_exposure_m1, m_cam_exposure_m2; uint32_t m_cam_exposure_iter = 0, m_n_steps_since_cam_update = 0, m_cam_exposure_n = 0;
    void ensure_cam_exposure(cudaStream_t s);
    // image mode
    DeviceBuffer m_image; int m_image_w = 0, m_image_h = 0; DeviceBuffer m_img_pos, m_img_target, m_img_out, m_img_dl, m_img_loss;
    // SDF mode (scratch shared with the image mode)
    DeviceBuffer m_sdf_pos, m_sdf_dist; uint32_t m_sdf_size = 0; bool m_sdf_step_skipped = false; int m_tcnn_loss = TLOSS_L2;
    // render scratch
    DeviceBuffer m_r_rgba[3], m_r_depth[3], m_r_payload[3], m_r_input, m_r_output, m_r_counters, m_r_scan, m_frame, m_frame_depth, m_accum, m_surface, m_r_coarse;
    TraceCtl* m_trace_host = nullptr; cudaEvent_t m_ev_trace[8] = {};   // host-mapped copies of the tracer's loop control + one event per ring slot
    // one sample of one row band through the wavefront tracer into m_frame / m_frame_depth (render_nerf, testbed_nerf.cu:2752-2871)
    void trace_sample(cudaStream_t s, const RenderCamera& cam, int sample, float* frame, float* fdepth);
};

}  // namespace ngp
