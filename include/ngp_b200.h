/* ngp_b200 — C ABI of the B200-native (sm_100a) instant-ngp hot path.
 *
 * Drop-in boundary for the path SURVEY.md §8 names: hash-grid encoding, fully fused 64-wide MLPs, the NerfNetwork
 * composition, fused Adam, and the Testbed NeRF / image training step + inference render.  Plain pointers and sizes
 * only; every `const void* dev_*` / `void* dev_*` argument is a DEVICE pointer, `stream` is a cudaStream_t passed as
 * void* (NULL = default stream).  All functions return 0 on success, non-zero on failure; ngp_last_error() returns
 * the message of the last failure on the calling thread (the reference throws std::runtime_error:
 * CUDA_CHECK_THROW, main.cu:195-209 — the Python binding re-raises it as RuntimeError like pybind11 does).
 * There is NO CPU fallback: without a CUDA device every compute entry point fails.
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference repo root).
 */
#ifndef NGP_B200_H
#define NGP_B200_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NGP_B200_ABI_VERSION 1
#define NGP_BATCH_GRANULARITY 256u /* tcnn::batch_size_granularity as used at testbed_nerf.cu:2601-2603,3606,3928 */

const char* ngp_last_error(void);
int ngp_abi_version(void);
/* number of this library's kernels launched so far on this process (bench.py's gpu_launches) */
uint64_t ngp_kernel_launch_count(void);
/* opt-in per-kernel CUDA-event timing on the launching stream (bench.py roofline leg); report = JSON object
 * {"<kernel group>": {"count": launches, "ms": total}} written to buf (NUL terminated) */
void ngp_profile_enable(int on);
int ngp_profile_report(char* buf, uint64_t buf_size, int reset);
/* the same records as an absolute timeline, [["kernel group", start_us, duration_us], ...] relative to the first record since the last
 * reset (both streams: shows how the ray prefetch overlaps the training stream); call before ngp_profile_report(.., reset = 1) */
int ngp_profile_timeline(char* buf, uint64_t buf_size);

/* Roofline denominators of the hash-grid kernels, measured on the current device (no reference
 * counterpart — the reference publishes no roofline): random 4-byte gathers per second with 32 loads in flight per
 * thread over a buffer of `buffer_bytes` (the table size: L2-resident when it fits), and scattered
 * red.global.add.f16x2 lanes per second (`vec` = 4-byte entries per lane: 1, 2 or 4).  Median of `reps` launches. */
int ngp_microbench_l2_gather(void* stream, uint64_t buffer_bytes, uint32_t reps, double* gathers_per_s);
int ngp_microbench_l2_red(void* stream, uint64_t buffer_bytes, int32_t vec, uint32_t reps, double* lanes_per_s);
/* Host-side test ngp_testbed_set_training_image applies to an RGBA8 view before it keeps a prefetched walk: does the view contain the
 * mask colour 0x00FF00FF of read_rgba (common_device.cuh:885-913)?  Exposed for tests (the scan reads eight interleaved streams: as one
 * linear stream it made the upload path host-bound).  No device needed. */
int ngp_host_has_mask_pixels(const void* rgba8_pixels, uint64_t n_pixels, int32_t* has_mask);

/* activation layouts for encoder outputs / MLP inputs */
enum { NGP_LAYOUT_ROWMAJOR = 0, /* [n][C] fp16 ("AoS", tcnn::RM with one sample per row) */
       NGP_LAYOUT_BLK8 = 1      /* 8x8 fp16 core matrices, the tcgen05 no-swizzle operand layout */ };

/* ------------------------------------------------------------------------------------------------------------
 * Encoding  —  replaces tcnn::create_encoding<T>(n_dims_to_encode, json, alignment) + Encoding::{forward,backward}
 *   reference call sites: include/neural-graphics-primitives/nerf_network.h:93-95,143-148,201-207,324-333,345-350;
 *   src/testbed.cu:3961-4005 (config derivation), :4101
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct ngp_encoding* ngp_encoding_t;
/* json: the "encoding" object of a network config (otype HashGrid, n_levels, n_features_per_level,
 * log2_hashmap_size, base_resolution, per_level_scale) */
int ngp_encoding_create(const char* json, uint32_t n_dims_to_encode, ngp_encoding_t* out);
void ngp_encoding_destroy(ngp_encoding_t e);
uint64_t ngp_encoding_n_params(ngp_encoding_t e);               /* Encoding::n_params() */
uint32_t ngp_encoding_padded_output_width(ngp_encoding_t e);    /* Encoding::padded_output_width(), alignment 16 */
/* level geometry: offsets[L+1] (entries), scales[L], resolutions[L]; any pointer may be NULL. returns n_levels */
uint32_t ngp_encoding_level_table(ngp_encoding_t e, uint32_t* offsets, float* scales, uint32_t* resolutions);
/* Encoding::initialize_params: U(-1e-4,1e-4), counter-based pcg32(seed) */
int ngp_encoding_initialize_params(ngp_encoding_t e, void* stream, uint64_t seed, void* dev_params_fp32);
/* Encoding::inference_mixed_precision / forward: x fp32 [n][x_stride], y fp16 [n][padded_width] in y_layout */
int ngp_encoding_forward(ngp_encoding_t e, void* stream, const void* dev_params_fp16, const void* dev_x, uint32_t x_stride,
                         uint32_t n, void* dev_y, int y_layout);
/* Encoding::backward, param gradients: dev_grad_fp16 (table shaped) += scatter(w * dy)  (EGradientMode::Accumulate;
 * zero it first for Overwrite) */
int ngp_encoding_backward(ngp_encoding_t e, void* stream, const void* dev_x, uint32_t x_stride, uint32_t n, const void* dev_dy,
                          int dy_layout, void* dev_grad_fp16);
/* GridEncoding::set_max_level_gpu (tiny-cuda-nn; call sites testbed_nerf.cu:3996,4004,4130): a device array with one float per row for the
 * forward / backward calls that follow.  Row i encodes the levels l < max_level[i] * n_levels + 1e-3 (the others are zeros) and scatters
 * gradients to the levels l <= max_level[i] * n_levels + 1e-3.  NULL (the default) = every level. */
int ngp_encoding_set_max_level_gpu(ngp_encoding_t e, const float* dev_max_level);
/* Encoding::backward, input gradients (testbed_nerf.cu:2615-2617,4071-4078): dx fp32 [n][n_dims] */
int ngp_encoding_backward_input(ngp_encoding_t e, void* stream, const void* dev_params_fp16, const void* dev_x, uint32_t x_stride,
                                uint32_t n, const void* dev_dy, int dy_layout, void* dev_dx);
/* parity hook: level-local corner indices uint32 [n][2^D] and fp32 weights [n][2^D] of one level */
int ngp_encoding_corners(ngp_encoding_t e, void* stream, const void* dev_x, uint32_t x_stride, uint32_t n, uint32_t level,
                         void* dev_idx, void* dev_w);

/* ------------------------------------------------------------------------------------------------------------
 * NerfNetwork  —  replaces ngp::NerfNetwork<T> (include/neural-graphics-primitives/nerf_network.h:76-578)
 * json: full network config with "encoding", "network", "dir_encoding", "rgb_network" (configs/nerf/base.json)
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct ngp_nerf_network* ngp_nerf_network_t;
int ngp_nerf_network_create(const char* json, ngp_nerf_network_t* out);
void ngp_nerf_network_destroy(ngp_nerf_network_t n);
uint64_t ngp_nerf_network_n_params(ngp_nerf_network_t n);                   /* nerf_network.h:459-461 */
/* flat layout (nerf_network.h:430-443): offsets of density MLP, rgb MLP, position encoding; n_matrix = MLP params */
int ngp_nerf_network_layout(ngp_nerf_network_t n, uint64_t* off_density, uint64_t* off_rgb, uint64_t* off_encoding, uint64_t* n_matrix_params);
int ngp_nerf_network_initialize_params(ngp_nerf_network_t n, void* stream, uint64_t seed, void* dev_params_fp32);  /* :445-457 */
/* inference_mixed_precision_impl (:116-173): coords fp32 [n][7] = (pos3, dt, dir3); out fp16 [n][4] = (r,g,b,sigma) raw.
 * n must be a multiple of 128 (callers round to NGP_BATCH_GRANULARITY like the reference). */
int ngp_nerf_network_inference(ngp_nerf_network_t n, void* stream, const void* dev_params_fp16, const void* dev_coords, uint32_t count, void* dev_out);
/* density() (:337-353): pos fp32 [n][pos_stride]; out fp16 [n] */
int ngp_nerf_network_density(ngp_nerf_network_t n, void* stream, const void* dev_params_fp16, const void* dev_pos, uint32_t pos_stride, uint32_t count, void* dev_out);
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same computation, same kernels and same arguments as
 * ngp_nerf_network_inference, but the call ALLOCATES NOTHING: the encoder-output scratch that the regular entry point keeps in a
 * grow-on-demand cudaMalloc'ed buffer lives in a file-scope static __device__ array (csrc/noalloc.cu) of
 * NGP_NOALLOC_MAX_SAMPLES x 32 fp16 = 1 MiB, held by every GPU that loads the library until the process exits.
 * count must be a multiple of 128 and <= NGP_NOALLOC_MAX_SAMPLES (larger counts fail with a clear error: split the batch).
 * Every call shares that array (it keeps the last call's contents; the call overwrites all it reads, nothing needs clearing):
 * calls that overlap on the GPU — different streams or host threads — RACE on it and must be serialised by the caller.
 * Per-kernel profiling (ngp_profile_enable) must be off for the no-allocation guarantee. */
#define NGP_NOALLOC_MAX_SAMPLES 16384u
uint32_t ngp_noalloc_max_samples(void);
int ngp_nerf_network_inference_noalloc(ngp_nerf_network_t n, void* stream, const void* dev_params_fp16, const void* dev_coords, uint32_t count,
                                       void* dev_out);
/* forward_impl + backward_impl with EGradientMode::Overwrite (:179-335): dL_dout fp16 [n][4]; grads fp16, param shaped */
int ngp_nerf_network_forward_backward(ngp_nerf_network_t n, void* stream, const void* dev_params_fp16, const void* dev_coords, uint32_t count,
                                      const void* dev_dL_dout, void* dev_grads_fp16);
/* the same with NerfNetwork::backward's dL_dinput (nerf_network.h:262,282-330; what train_nerf_step asks for when cameras are being optimised,
 * testbed_nerf.cu:4071-4078): dev_dL_dcoords fp32 [n][7] receives the gradient w.r.t. the warped position (cols 0..2, through the hash grid) and
 * the warped direction (cols 4..6, through the spherical harmonics); col 3 (dt) is not a network input and is not written */
int ngp_nerf_network_forward_backward_input(ngp_nerf_network_t n, void* stream, const void* dev_params_fp16, const void* dev_coords, uint32_t count,
                                            const void* dev_dL_dout, void* dev_grads_fp16, void* dev_dL_dcoords);

/* ------------------------------------------------------------------------------------------------------------
 * NetworkWithInputEncoding — replaces tcnn::NetworkWithInputEncoding<T> (src/testbed.cu:4101-4113; image/SDF modes)
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct ngp_encoded_mlp* ngp_encoded_mlp_t;
int ngp_encoded_mlp_create(const char* json, uint32_t n_input_dims, uint32_t n_output_dims, ngp_encoded_mlp_t* out);
void ngp_encoded_mlp_destroy(ngp_encoded_mlp_t m);
uint64_t ngp_encoded_mlp_n_params(ngp_encoded_mlp_t m);
int ngp_encoded_mlp_layout(ngp_encoded_mlp_t m, uint64_t* off_mlp, uint64_t* off_encoding, uint64_t* n_matrix_params);
int ngp_encoded_mlp_initialize_params(ngp_encoded_mlp_t m, void* stream, uint64_t seed, void* dev_params_fp32);
/* x fp32 [n][x_stride]; out fp16 [n][16] (first n_output_dims meaningful) */
int ngp_encoded_mlp_inference(ngp_encoded_mlp_t m, void* stream, const void* dev_params_fp16, const void* dev_x, uint32_t x_stride, uint32_t count, void* dev_out);
int ngp_encoded_mlp_forward_backward(ngp_encoded_mlp_t m, void* stream, const void* dev_params_fp16, const void* dev_x, uint32_t x_stride, uint32_t count,
                                     const void* dev_dL_dout /* fp16 [n][16] */, void* dev_grads_fp16);

/* ------------------------------------------------------------------------------------------------------------
 * Optimizer step — replaces tcnn::Trainer::optimizer_step -> Ema(ExponentialDecay(Adam)) (src/testbed_nerf.cu:3678,
 * src/testbed_image.cu:283).  Stateless form: the caller owns all buffers.
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct ngp_adam_hyper {
    float learning_rate, beta1, beta2, epsilon, l2_reg, loss_scale;
    uint64_t n_matrix_params;
    int32_t optimize_matrix_params, optimize_non_matrix_params;
    float ema_decay;      /* 0 = no EMA */
    uint32_t ema_step;    /* 1-based */
} ngp_adam_hyper;
int ngp_adam_step(void* stream, const ngp_adam_hyper* h, uint64_t n, void* dev_master_fp32, void* dev_params_fp16, const void* dev_grads_fp16,
                  void* dev_m_fp32, void* dev_v_fp32, void* dev_steps_u32, void* dev_ema_fp32 /* nullable */, void* dev_ema_fp16 /* nullable */);


/* ------------------------------------------------------------------------------------------------------------
 * Testbed — replaces the training / inference slice of ngp::Testbed as exposed by the pyngp module
 * (src/python_api.cu:266-538; include/neural-graphics-primitives/testbed.h:64-516).
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct ngp_testbed* ngp_testbed_t;
enum { NGP_MODE_NERF = 0, NGP_MODE_SDF = 1, NGP_MODE_IMAGE = 2 };             /* ETestbedMode (python_api.cu:271-277) */
enum { NGP_IMAGE_RGBA8 = 0, NGP_IMAGE_RGBA32F = 1 };                         /* EImageDataType Byte / Float */
enum { NGP_LENS_PERSPECTIVE = 0, NGP_LENS_OPENCV = 1 };                      /* ELensMode (common.h:243-250) */
enum { NGP_ACT_NONE = 0, NGP_ACT_RELU = 1, NGP_ACT_LOGISTIC = 2, NGP_ACT_EXPONENTIAL = 3 };   /* ENerfActivation */
enum { NGP_LOSS_L2 = 0, NGP_LOSS_L1 = 1, NGP_LOSS_MAPE = 2, NGP_LOSS_SMAPE = 3, NGP_LOSS_HUBER = 4, NGP_LOSS_LOGL1 = 5, NGP_LOSS_RELATIVE_L2 = 6 };

/* one training view in HOST memory: what NerfDataset::set_training_image + metadata/xforms hold
 * (nerf_loader.h:31-46,110-171; python_api.cu:49-67 set_image, :583-605 set_camera_intrinsics/extrinsics).
 * xform = the NGP-convention camera-to-world mat4x3, column major (x axis, y axis, z axis, position). */
typedef struct ngp_training_image {
    const void* pixels; int32_t data_type; int32_t res[2]; float focal_length[2]; float principal_point[2];
    int32_t lens_mode; float lens_params[4]; float xform[12];
} ngp_training_image;

typedef struct ngp_train_stats {   /* Testbed.training_step / .loss / nerf.training.counters_rgb */
    uint32_t training_step; float loss; uint32_t rays_per_batch, measured_batch_size, measured_batch_size_before_compaction, n_rays_total;
    float learning_rate; float density_grid_mean;
} ngp_train_stats;

int ngp_testbed_create(int mode, ngp_testbed_t* out);                                   /* Testbed(mode) python_api.cu:388 */
void ngp_testbed_destroy(ngp_testbed_t t);
int ngp_testbed_reload_network_from_json(ngp_testbed_t t, const char* json);            /* python_api.cu:402-407 */
/* reload_network_from_json(json, config_base_path) (python_api.cu:402-407, testbed.cu:311-316): a config with a "parent" key is an RFC 7386
 * merge patch over the file <directory of config_base_path>/<parent>, resolved recursively (merge_parent_network_config, testbed.cu:95-106);
 * name a file, or a directory with a trailing slash */
int ngp_testbed_reload_network_from_json_base(ngp_testbed_t t, const char* json, const char* config_base_path);
/* reload_network_from_file(path = "") (python_api.cu:400-401, testbed.cu:228-309): `path` if it exists, else <root_dir>/configs/<nerf|sdf|image>/<path>
 * for a relative one (find_network_config); JSON with comments; "parent" configs merged; then reset_network.  The path ARGUMENT is remembered
 * (m_network_config_path, "base.json" at first) and an empty / NULL path reloads it.  ngp_testbed_train on a testbed that has no network yet does
 * exactly that (testbed.cu:4297-4306).  When "base.json" is not found under root_dir the hyper-parameters of the reference's configs/<mode>/base.json
 * that are built into the library are used (ngp_builtin_base_config returns that JSON text; mode 0 Nerf, 1 Sdf, 2 Image).  Snapshot files go through
 * ngp_testbed_load_snapshot.  ngp_testbed_set_root_dir = Testbed.root_dir (python_api.cu:534-537). */
int ngp_testbed_reload_network_from_file(ngp_testbed_t t, const char* path);
int ngp_testbed_set_root_dir(ngp_testbed_t t, const char* root_dir);
int ngp_builtin_base_config(int32_t mode, const char** json);
/* load_training_data (python_api.cu:398) with in-memory views; uploads pixels host->device */
int ngp_testbed_set_nerf_dataset(ngp_testbed_t t, void* stream, uint32_t n_images, const ngp_training_image* images, int32_t aabb_scale);
/* load_training_data(path) for a NeRF scene (python_api.cu:398; load_nerf, src/nerf_loader.cu:259-726): the host side of the loader.
 * ngp_nerf_dataset_open parses one or more transforms.json files (paths separated by '\n') and does everything load_nerf derives
 * from them: natural-order frame sort, n_frames / sharpness culling, scene scale / offset / aabb / aabb_scale / up, per-frame
 * lens + principal point + NeRF -> NGP pose.  Decoding the image files is the caller's job (the reference uses stb_image, a
 * third-party dependency): for each frame i read ngp_nerf_dataset_image_path, decode, and hand the pixels ([h][w][4] uint8 sRGB
 * straight alpha = data_type 0, or float32 = 1) to ngp_nerf_dataset_set_image, which also resolves the focal length (the fov
 * forms need the resolution).  ngp_testbed_load_nerf_dataset then uploads the views (= set_nerf_dataset with the file's aabb_scale). */
typedef struct ngp_nerf_dataset* ngp_nerf_dataset_t;
int ngp_nerf_dataset_open(const char* json_paths, ngp_nerf_dataset_t* out);
void ngp_nerf_dataset_close(ngp_nerf_dataset_t d);
uint32_t ngp_nerf_dataset_n_images(ngp_nerf_dataset_t d);
int ngp_nerf_dataset_image_path(ngp_nerf_dataset_t d, uint32_t i, char* buf, uint64_t buf_size);
int ngp_nerf_dataset_set_image(ngp_nerf_dataset_t d, uint32_t i, const void* host_pixels, int32_t data_type, int32_t width, int32_t height);
/* depth supervision data (nerf_loader.cu:471-473,609-621,711): with "integer_depth_scale" in the scene json a frame may name a "depth_path", a
 * 16-bit single-channel image of the view's resolution.  ngp_nerf_dataset_depth_path gives the resolved path ("" = nothing to read for this frame);
 * the caller decodes it and hands the [h][w] uint16 pixels over (after the frame's ngp_nerf_dataset_set_image).  ngp_testbed_load_nerf_dataset
 * uploads value * integer_depth_scale * scene scale per pixel. */
int ngp_nerf_dataset_depth_path(ngp_nerf_dataset_t d, uint32_t i, char* buf, uint64_t buf_size);
int ngp_nerf_dataset_set_depth(ngp_nerf_dataset_t d, uint32_t i, const uint16_t* host_pixels, int32_t width, int32_t height);
/* scene-level results: aabb_scale, scale, offset[3], from_mitsuba */
int ngp_nerf_dataset_info(ngp_nerf_dataset_t d, int32_t* aabb_scale, float* scale, float* offset3, int32_t* from_mitsuba);
/* view i as it will be uploaded (pixels point into the dataset's own copy; focal length valid after set_image) */
int ngp_nerf_dataset_view(ngp_nerf_dataset_t d, uint32_t i, ngp_training_image* out);
/* the end pose ("transform_matrix_end", NGP convention; equal to the start pose when the frame has none) and the rolling-shutter coefficients
 * ("rolling_shutter": [A, B, C(, D)] of the scene or the frame) of frame i (nerf_loader.cu:190-202,648-649,676-680) */
int ngp_nerf_dataset_view_motion(ngp_nerf_dataset_t d, uint32_t i, float* xform_end12, float* rolling_shutter4);
int ngp_testbed_load_nerf_dataset(ngp_testbed_t t, void* stream, ngp_nerf_dataset_t d);
/* nerf.training.set_image(frame_idx, img) (python_api.cu:49-67,583): replace one view's pixels (same size and type as
 * given to set_nerf_dataset) from HOST memory.  The copy is ordered after everything already queued on `stream` and before the
 * next training step's loss kernel; RGBA8 views travel on an internal copy stream so that the upload runs beside the start of
 * that step.  The host buffer must stay valid until the next ngp_testbed_train* call has returned. */
int ngp_testbed_set_training_image(ngp_testbed_t t, void* stream, uint32_t frame_idx, const void* host_pixels);
/* nerf.training.set_camera_intrinsics(frame_idx, fx, fy, cx, cy, k1, k2, p1, p2) / set_camera_extrinsics(frame_idx, camera_to_world,
 * convert_to_ngp) (python_api.cu:583-605, testbed_nerf.cu:2873-2919): cx, cy in pixels when >= 0, a normalised principal point
 * when negative; all-zero distortion = Perspective; camera_to_world = mat4x3 column major, NeRF convention when convert_to_ngp.
 * An out-of-range frame index is ignored, as in the reference. */
int ngp_testbed_set_camera_intrinsics(ngp_testbed_t t, void* stream, int32_t frame_idx, float fx, float fy, float cx, float cy,
                                      float k1, float k2, float p1, float p2);
int ngp_testbed_set_camera_extrinsics(ngp_testbed_t t, void* stream, int32_t frame_idx, const float* camera_to_world, int32_t convert_to_ngp);
/* nerf.training.set_camera_extrinsics_rolling_shutter(frame_idx, camera_to_world_start, camera_to_world_end, rolling_shutter, convert_to_ngp)
 * (python_api.cu:606-617, testbed_nerf.cu:2896-2915): the pose at the start and at the end of the exposure and (A, B, C, D); every training ray
 * of the view gets the camera interpolated at pixel time A + B u + C v + D motionblur_time (get_xform_given_rolling_shutter, common_device.cuh:401-408:
 * position linearly, rotation by quaternion slerp).  The loader sets both from "transform_matrix_start" / "_end" and "rolling_shutter". */
int ngp_testbed_set_camera_extrinsics_rolling_shutter(ngp_testbed_t t, void* stream, int32_t frame_idx, const float* camera_to_world_start,
                                                      const float* camera_to_world_end, const float* rolling_shutter4, int32_t convert_to_ngp);
/* nerf.training.get_camera_extrinsics(frame_idx, convert_to_ngp = True) (python_api.cu:594-603, testbed_nerf.cu:2925-2932): the pose of view frame_idx
 * as training sees it - the given pose with the trained rotation / position offsets of optimize_extrinsics applied - as a column-major mat4x3,
 * converted back to the NeRF convention when convert_to_ngp is set (the flag is named as in the reference: it means "the matrix I want is in the
 * convention I would pass WITH convert_to_ngp") */
int ngp_testbed_get_camera_extrinsics(ngp_testbed_t t, int32_t frame_idx, int32_t convert_to_ngp, float* out12);
/* the depth half of NerfDataset::set_training_image (nerf_loader.cu:907-959; nerf.training.set_image(frame_idx, img, depth_img, depth_scale),
 * python_api.cu:49-67): HOST depth map of view frame_idx, [h][w] uint16 (depth_type 0) or float32 (1), kept on the device as value * depth_scale
 * (scene units along the view axis).  NULL or depth_scale == 0: a map of zeros (no supervision from this view); depth_scale < 0: the view's map is
 * dropped.  Used by the loss when the option depth_supervision_lambda > 0 (testbed_nerf.cu:1848-1856,1934-1939; depth_loss_type, L1 by default). */
int ngp_testbed_set_training_depth(ngp_testbed_t t, void* stream, int32_t frame_idx, const void* host_depth, int32_t depth_type, float depth_scale);
/* Testbed.reset(reset_density_grid) (python_api.cu: reset_network): fresh parameters, optimizer state and counters */
int ngp_testbed_reset(ngp_testbed_t t, int32_t reset_density_grid);
int ngp_testbed_set_image(ngp_testbed_t t, void* stream, const float* host_rgba, int32_t width, int32_t height);   /* image mode */
/* image mode, an 8-bit file as load_stbi_image keeps it (testbed_image.cu:404-418; load_stbi_gpu + from_rgba32, common_device.cu:47-76,
 * common_device.cuh:771-798): HOST [h][w][4] sRGB bytes with straight alpha become linear, alpha-premultiplied floats on the device */
int ngp_testbed_set_image_rgba8(ngp_testbed_t t, void* stream, const uint8_t* host_rgba8, int32_t width, int32_t height);
/* mode_from_scene (common.cu:144-159), what Testbed::load_training_data picks its mode with (testbed.cu:144-150): a directory or *.json ->
 * 0 Nerf, *.obj / *.stl -> 1 Sdf, *.nvdb -> 3 Volume, any other existing file -> 2 Image, a path that does not exist -> 4 None
 * (ETestbedMode's values; this library has testbeds for 0, 1 and 2) */
int ngp_mode_from_scene(const char* path, int32_t* mode);
/* the mode a snapshot file was saved in (testbed.cu:4950-4955: Testbed::load_snapshot switches to it before it reads anything else);
 * same values.  No device needed. */
int ngp_snapshot_mode(const char* path, int32_t* mode);
/* Sdf mode, Testbed::override_sdf_training_data (python_api.cu:69-99,437): n HOST (point xyz, signed distance) pairs in mesh
 * space; raw_aabb_min/max and mesh_scale are what load_mesh derives from the mesh (m_raw_aabb, m_sdf.mesh_scale) and move the
 * points into the unit cube (pass {0,0,0},{1,1,1},1 for data that is already there).  ngp_testbed_train then runs train_sdf
 * (testbed_sdf.cu:1289-1311): the first batch_size records of a permutation keyed by the training step, config loss (MAPE for
 * configs/sdf/base.json), optimizer.  As in the reference, a step is skipped when the data set is smaller than the batch. */
int ngp_testbed_override_sdf_training_data(ngp_testbed_t t, void* stream, const float* host_points, const float* host_distances, uint32_t n,
                                           const float* raw_aabb_min, const float* raw_aabb_max, float mesh_scale);
int ngp_testbed_train(ngp_testbed_t t, void* stream, uint32_t batch_size);              /* Testbed::train, python_api.cu:426 */
/* data-parallel split of train(): gradients only / optimizer + counters.  Between the two calls the caller all-reduces
 * (sum) the fp16 gradient buffer; measured_* = 0 keeps the local sample counts for the rays-per-batch control law. */
int ngp_testbed_train_begin(ngp_testbed_t t, void* stream, uint32_t batch_size);
int ngp_testbed_train_end(ngp_testbed_t t, void* stream, uint32_t measured_batch_size, uint32_t measured_batch_size_before_compaction);
/* local (measured, measured_before_compaction) of the step begun by train_begin; returns as soon as the loss kernel has
 * run, while the backward pass is still in flight, so that the host can exchange the counts under it
 * (NerfCounters, testbed.h:433-447) */
int ngp_testbed_pending_counters(ngp_testbed_t t, uint32_t* measured_batch_size, uint32_t* measured_batch_size_before_compaction);
int ngp_testbed_set_shard(ngp_testbed_t t, uint32_t rank, uint32_t world_size);
/* Data-parallel step inside the library (ngp_comm_init / ngp_allreduce_gradients).  Rank 0 obtains an NCCL unique id
 * (128 bytes), the host distributes it, every rank calls ngp_testbed_comm_init (collective).  From then on ngp_testbed_train
 * shards the rays by rank, all-reduces the two sample counters under the backward pass and then the whole fp16 gradient vector
 * (layout nerf_network.h:430-443) in ONE collective behind the backward pass on a communication stream (cutting it into
 * hash-level slices that overlap the scatter was measured slower at this message size; NGP_DP_SLICES
 * re-enables it for experiments); every rank then runs the identical optimizer step — or, with "dp_sharded_optimizer" = 1,
 * reduce-scatter -> Adam on this rank's 1/world of the entries -> all-gather of the fp16 parameters.  NCCL is bound at run
 * time: nccl_lib_path names libnccl.so.2 (NULL: the copy already loaded in the process, $NGP_NCCL_LIB, or the system one). */
int ngp_comm_unique_id(void* out_128_bytes, const char* nccl_lib_path);
int ngp_testbed_comm_init(ngp_testbed_t t, const void* id_128_bytes, uint32_t rank, uint32_t world_size, const char* nccl_lib_path);
/* Sharded optimizer ("dp_sharded_optimizer" = 1, data parallel only; no reference counterpart — the reference trains on one GPU,
 * README.md:250-252): the EMA / inference parameters of the other ranks' slices are fetched with one in-place ncclAllGather.
 * COLLECTIVE: every rank calls it before any rank renders or saves a snapshot; a no-op otherwise. */
int ngp_testbed_sync_inference_params(ngp_testbed_t t, void* stream);
int ngp_testbed_stats(ngp_testbed_t t, ngp_train_stats* out);
uint64_t ngp_testbed_n_params(ngp_testbed_t t);                                         /* Testbed::n_params python_api.cu:440 */
uint64_t ngp_testbed_n_encoding_params(ngp_testbed_t t);                                /* Testbed::n_encoding_params python_api.cu:445 */
/* Testbed::save_snapshot(path, include_optimizer_state, compress) / load_snapshot(path) (python_api.cu:446-447; testbed.cu:4873-5057):
 * MessagePack of the network config + "snapshot" {n_params, params_type "__half", params_binary (fp16 inference parameters),
 * density_grid_size 128, density_grid_binary (fp16), nerf{aabb_scale, rgb{rays_per_batch, measured_batch_size, ...}, dataset{...}},
 * training_step, loss, aabb, render_aabb, background_color, version 1, mode, [optimizer]}; a path ending in ".ingp" is wrapped in a
 * zlib stream (compress = 0: stored), anything else (".msgpack") is written bare.  load_snapshot rebuilds the network from the
 * embedded config, so the testbed needs no prior reload_network call; its mode must be the snapshot's. */
/* host-only helpers of the snapshot codec (no device needed): a network config as JSON text (comments allowed, as the reference
 * parses it) -> MessagePack, and decode + re-encode of a MessagePack buffer (what load_snapshot followed by save_snapshot does to the
 * embedded config).  *out_size receives the encoded size; status 1 when `capacity` is too small or the input is malformed. */
int ngp_json_to_msgpack(const char* json, void* out, uint64_t capacity, uint64_t* out_size);
int ngp_msgpack_roundtrip(const void* in, uint64_t in_size, void* out, uint64_t capacity, uint64_t* out_size);
int ngp_testbed_save_snapshot(ngp_testbed_t t, void* stream, const char* path, int32_t include_optimizer_state, int32_t compress);
int ngp_testbed_load_snapshot(ngp_testbed_t t, void* stream, const char* path);
/* device pointers of the trainer's flat buffers (layout nerf_network.h:430-443): 0 master fp32, 1 params fp16,
 * 2 inference params fp16 (EMA), 3 gradients fp16, 4 / 5 Adam first / second moments fp32, 6 per-entry step counts u32,
 * 7 EMA fp32 (NULL without an Ema optimizer) */
int ngp_testbed_param_buffer(ngp_testbed_t t, int which, void** dev_ptr, uint64_t* n_elements);
/* occupancy grid: 0 density grid fp32 [128^3*(max_cascade+1)], 1 bitfield u8 [128^3*8/8] */
int ngp_testbed_density_grid(ngp_testbed_t t, int which, void** dev_ptr, uint64_t* n_bytes);
/* Testbed::render (python_api.cu:127-205) headless: linear premultiplied RGBA float32 rows [y0,y1) of a w x h image,
 * averaged over spp; camera = mat4x3 column major; out is a HOST buffer unless out_is_device */
int ngp_testbed_render(ngp_testbed_t t, void* stream, const float* camera_matrix, const float* focal_length, const float* screen_center,
                       int32_t width, int32_t height, int32_t spp, int32_t y0, int32_t y1, float* rgba_out, int32_t out_is_device, float* depth_out);
/* Testbed::render_to_cpu (python_api.cu:127-205 -> render_frame testbed.cu:4640-4760) headless, with the testbed's OWN view state
 * (camera_matrix, fov / relative_focal_length, fov_axis, zoom, screen_center, render lens): spp samples are accumulated in "color_space"
 * (accumulate_kernel, render_buffer.cu:232-260), then composited over "background_color" with "exposure", "tonemap_curve" and the output
 * colour space (sRGB unless `linear`; tonemap_kernel, render_buffer.cu:533-565), then the ground-truth overlay of the selected training
 * view when "render_ground_truth" is set (overlay_image_kernel, render_buffer.cu:348-418).  float32 RGBA rows [y0,y1) of a w x h frame
 * into a HOST buffer unless out_is_device.  Camera-path arguments of the reference (start_t / end_t / fps / shutter_fraction) belong to
 * the GUI's camera path, which is out of scope: the binding rejects them. */
int ngp_testbed_render_frame(ngp_testbed_t t, void* stream, int32_t width, int32_t height, int32_t spp, int32_t linear, int32_t y0, int32_t y1,
                             void* rgba_out, int32_t out_is_device);
/* Testbed::reset_camera (testbed.cu:858-884), set_nerf_camera_matrix (:789-791; mat4x3 column major in the NeRF convention),
 * set_camera_to_training_view (:841-856: pose, focal length, principal point and lens of training view i become the view state) */
int ngp_testbed_reset_camera(ngp_testbed_t t);
int ngp_testbed_set_nerf_camera_matrix(ngp_testbed_t t, const float* nerf_camera_to_world_12);
int ngp_testbed_set_camera_to_training_view(ngp_testbed_t t, int32_t trainview);
/* vector-valued members (python_api.cu:452-537): "camera_matrix" (12, mat4x3 column major, NGP convention), "screen_center" (2),
 * "relative_focal_length" (2), "up_dir" (3), "background_color" (4, sRGB RGBA), "aabb" / "render_aabb" (6: min xyz, max xyz) */
int ngp_testbed_set_vector(ngp_testbed_t t, const char* name, const float* values, uint32_t count);
int ngp_testbed_get_vector(ngp_testbed_t t, const char* name, float* values, uint32_t count);
/* nerf.render_lens / render_with_lens_distortion (set_camera_to_training_view): lens model applied to the rays of
 * ngp_testbed_render; lens_mode 0 Perspective, 1 OpenCV (k1, k2, p1, p2) */
int ngp_testbed_set_render_lens(ngp_testbed_t t, int32_t lens_mode, const float* lens_params4);
/* metadata of training view i as loaded (nerf.training.dataset.metadata[i] + xforms[i]); pixels is the DEVICE copy */
int ngp_testbed_training_view(ngp_testbed_t t, uint32_t i, ngp_training_image* out);
/* nerf.training.dataset.transforms[i] / metadata[i].rolling_shutter (python_api.cu:571-581,596): the start and end pose of view i AS GIVEN (NGP convention,
 * column-major mat4x3; the poses training uses carry the trained offsets on top: ngp_testbed_get_camera_extrinsics) and its rolling-shutter
 * coefficients; any pointer may be NULL.  The data set's scale / offset / aabb_scale: options "dataset_scale", "dataset_aabb_scale", vector "dataset_offset". */
int ngp_testbed_training_view_motion(ngp_testbed_t t, uint32_t i, float* xform_start12, float* xform_end12, float* rolling_shutter4);
uint32_t ngp_testbed_n_training_views(ngp_testbed_t t);
int ngp_testbed_compute_image_mse(ngp_testbed_t t, void* stream, int32_t quantize_to_byte, float* mse_out);   /* python_api.cu:444 */
/* runtime knobs (pyngp def_readwrite members, python_api.cu:452-537,611-695); value semantics per name:
 * "random_bg_color","linear_colors","shall_train_encoding","shall_train_network" (bool),
 * "train_snap_to_pixel_centers" (bool, default 1: nerf.training.snap_to_pixel_centers, training rays),
 * "snap_to_pixel_centers" (bool, default 0: Testbed::m_snap_to_pixel_centers, render rays: pixel offset of sample 0 for every sample),
 * "near_distance","density_grid_decay","render_min_transmittance","cone_angle_constant" (float),
 * "loss_type","rgb_activation","density_activation" (enum), "seed" (integer), "image_snap_to_pixel_centers",
 * "image_linear_colors","image_stratified" (bool), "background_r/g/b/a" (float),
 * view + output stage: "fov" (degrees along fov_axis), "fov_axis" (0|1), "zoom", "scale", "exposure" (stops), "color_space"
 * (0 Linear | 1 SRGB | 2 VisPosNeg), "tonemap_curve" (0 Identity | 1 ACES | 2 Hable | 3 Reinhard), "render_mode" (ERenderMode: 0 AO,
 * 1 Shade, 3 Positions, 4 Depth, 6 Cost, 9 EncodingVis; 2 Normals is rejected), "render_ground_truth" (bool), "ground_truth_alpha",
 * "training_view", "render_with_lens_distortion" (bool), "render_near_distance",
 * "render_row_stride" (default 1; both render entry points then produce rows y0, y0 + stride, ... < y1: interleaved tile sharding),
 * "sampler_lanes_per_ray" (0 automatic | 1 | 2 | 4 | 8 | 32: lanes that cooperate on one training ray; results are identical),
 * "prefetch_rays" (bool, default 1: the next step's rays are generated on a second stream beside this step's backward pass
 * and optimizer whenever the next step does not refresh the occupancy grid first; results are identical),
 * "device_control_law" (bool, default 1: the rays_per_batch control law is ALSO evaluated on the device so that the next step's rays are
 * generated without waiting for the host to read this step's counters; results are identical),
 * "optimize_exposure" (bool, default 0), "exposure_l2_reg" (float, default 0), "n_steps_between_cam_updates" (default 16):
 * nerf.training.optimize_exposure (testbed_nerf.cu:1794,1969-1983,3831-3858) — per-image exposure trained on the device; debug buffers
 * "cam_exposure" / "cam_exposure_gradient" ([n_images][3] fp32) expose the values,
 * "walk_ahead" (bool, default 0 — measured slower, see csrc/testbed.h; scenes with cascades / a cone angle only: the counting walk of step N+2 is started during step N+1 with the
 * ray count predicted to stay the same, and redone when the control law says otherwise; results are identical; "walk_ahead_hits" /
 * "walk_ahead_misses" count the outcomes, read-only),
 * "dp_sharded_optimizer" (bool, default 0; data parallel: reduce-scatter -> Adam on 1/world of the entries -> all-gather) */
int ngp_testbed_set_option(ngp_testbed_t t, const char* name, double value);
/* parity hook: device pointer + capacity of an internal per-step buffer ("coords","numsteps","ray_indices","rays",
 * "counters","mlp_out","dloss","coords_compacted","loss","grid_pos","grid_idx","grid_out","mean_density",...) */
int ngp_testbed_debug_buffer(ngp_testbed_t t, const char* name, void** dev_ptr, uint64_t* n_bytes);
int ngp_testbed_get_option(ngp_testbed_t t, const char* name, double* value);


/* ------------------------------------------------------------------------------------------------------------
 * Synthetic dataset (BASELINE.json: 100 posed 800x800 views of a procedural analytic density field; no reference
 * counterpart — the reference loads transforms.json + images, src/nerf_loader.cu).  Data synthesis only.
 * ------------------------------------------------------------------------------------------------------------ */
/* camera `view` of `n_views` on an upper-hemisphere Fibonacci lattice (NGP convention mat4x3, host side) */
int ngp_synth_camera(uint32_t view, uint32_t n_views, float radius, float* xform_out);
/* ground-truth RGBA8 (straight-alpha sRGB) view into DEVICE memory [h][w] */
int ngp_synth_render_view(void* stream, uint32_t view, uint32_t n_views, float radius, float focal_px, int32_t width, int32_t height, void* dev_rgba8);

#ifdef __cplusplus
}
#endif
#endif /* NGP_B200_H */
