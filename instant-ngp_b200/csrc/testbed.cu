// Host driver (see testbed.h).  Control flow mirrors Testbed::train / train_nerf / train_nerf_step /
// update_density_grid_nerf / train_image of the reference; the kernels are the sm_100a ones of this library.
#include "testbed.h"
#include <fstream>
#include <sstream>
#include <sys/stat.h>
#include <algorithm>
#include <cmath>
#include <cstring>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace ngp {

static int tcnn_loss_from_string(const std::string& s) {   // tcnn::create_loss otypes used by the image / SDF configs
    const std::string l = to_lower(s);
    if (l == "l2") return TLOSS_L2; if (l == "relativel2") return TLOSS_RELATIVE_L2; if (l == "l1") return TLOSS_L1;
    if (l == "mape") return TLOSS_MAPE; if (l == "smape") return TLOSS_SMAPE;
    throw Error("Unknown loss type '" + s + "'");
}

static int loss_type_from_string(const std::string& s) {   // string_to_loss_type (testbed.cu:3872-3891)
    const std::string l = to_lower(s);
    if (l == "l2") return LOSS_L2; if (l == "l1") return LOSS_L1; if (l == "mape") return LOSS_MAPE; if (l == "smape") return LOSS_SMAPE;
    if (l == "huber" || l == "smoothl1") return LOSS_HUBER; if (l == "logl1") return LOSS_LOGL1; if (l == "relativel2") return LOSS_RELL2;
    throw Error("Unknown loss type '" + s + "'");
}

// RGBA8 pixels equal to 0x00FF00FF are "masked away" (read_rgba, common_device.cuh:885-913): the training walk skips rays that
// land on them, i.e. the walk depends on the image CONTENT only through this one value, and set_training_image needs to know — on the
// host and at once, before the next step may adopt a prefetched walk — whether a new view has any.  The scan reads the 2.56 MB of an
// 800x800 view from DRAM; as one linear stream that took 0.25-0.27 ms on the B200 box's host and made the upload path of bench.py
// HOST-bound (0.573 ms per step against 0.527 with the scan skipped).  Eight interleaved streams, a cache line each per iteration, keep
// more misses in flight: 0.15 ms (tools/scan_bench.cpp, profiles/r02_host_scan.txt).  A pool of sleeping worker threads was
// measured too: 0.07 ms in a tight loop, but with 0.5 ms between calls the wake-ups cost more than the scan (e2e 462 -> 325 M samples/s).
static bool has_mask_pixels(const void* pixels, size_t n_pixels) {
    const uint32_t* p = static_cast<const uint32_t*>(pixels);
    uint32_t any = 0;
    size_t done = 0;
#if defined(__SSE2__)
    constexpr size_t S = 8;
    const size_t c = (n_pixels / S) & ~(size_t)15;     // pixels per stream, whole cache lines
    __m128i acc = _mm_setzero_si128();
    const __m128i key = _mm_set1_epi32(0x00FF00FF);
    for (size_t i = 0; i < c; i += 16) {
        for (size_t s = 0; s < S; ++s) {
            const __m128i* q = reinterpret_cast<const __m128i*>(p + s * c + i);
            acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q), key)); acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 1), key));
            acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 2), key)); acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 3), key));
        }
    }
    any = _mm_movemask_epi8(acc) != 0;
    done = c * S;
#endif
    for (size_t i = done; i < n_pixels; ++i) any |= (uint32_t)(p[i] == 0x00FF00FFu);
    return any != 0;
}

bool host_has_mask_pixels(const void* rgba8_pixels, size_t n_pixels) { return has_mask_pixels(rgba8_pixels, n_pixels); }

Testbed::Testbed(int mode) : m_mode(mode), m_rng(1337), m_density_grid_rng(1337) {
    NGP_CHECK(mode == MODE_TB_NERF || mode == MODE_TB_IMAGE || mode == MODE_TB_SDF, "Testbed: unknown mode (Nerf = 0, Sdf = 1, Image = 2)");
    reset_camera();
}

Testbed::~Testbed() {
    if (m_aux_stream) { cudaStreamSynchronize(m_aux_stream); cudaStreamDestroy(m_aux_stream); }
    if (m_ev_counters) cudaEventDestroy(m_ev_counters);
    if (m_ev_prefetch) cudaEventDestroy(m_ev_prefetch);
    if (m_ev_after_mlp) cudaEventDestroy(m_ev_after_mlp);
    if (m_ev_after_loss) cudaEventDestroy(m_ev_after_loss);
    if (m_ev_walk_ahead) cudaEventDestroy(m_ev_walk_ahead);
    if (m_trace_host) { cudaFreeHost(m_trace_host); for (cudaEvent_t e : m_ev_trace) if (e) cudaEventDestroy(e); }
    if (m_comm_stream) { cudaStreamSynchronize(m_comm_stream); cudaStreamDestroy(m_comm_stream); }
    if (m_copy_stream) { cudaStreamSynchronize(m_copy_stream); cudaStreamDestroy(m_copy_stream); }
    for (cudaEvent_t e : {m_ev_copy_gate, m_ev_copy_done}) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : {m_ev_slice, m_ev_comm_done, m_ev_dp_counters}) if (e) cudaEventDestroy(e);
    m_comm.reset();
    if (m_host_counters) cudaFreeHost(m_host_counters);
    for (void* p : m_image_buffers) cudaFree(p);
    for (float* p : m_depth_buffers) if (p) cudaFree(p);
}

static std::string read_text_file(const std::string& path, const char* what) {
    std::ifstream f(path, std::ios::in | std::ios::binary);
    if (!f) throw Error(std::string(what) + " '" + path + "' does not exist.");
    std::stringstream ss; ss << f.rdbuf();
    return ss.str();
}
static bool path_exists(const std::string& p) { struct stat st; return !p.empty() && stat(p.c_str(), &st) == 0; }
static std::string join_path(const std::string& dir, const std::string& name) { return dir.empty() ? name : (dir.back() == '/' ? dir + name : dir + "/" + name); }
static std::string dir_of(const std::string& p) { const size_t k = p.find_last_of('/'); return k == std::string::npos ? std::string() : (k == 0 ? std::string("/") : p.substr(0, k)); }

// merge_parent_network_config (testbed.cu:95-106): a config names its "parent" relative to its own directory; the child is an RFC 7386
// merge patch on top of the (recursively resolved) parent
static Json merge_parent_network_config(const Json& child, const std::string& child_path, int depth = 0) {
    if (!child.contains("parent")) return child;
    NGP_CHECK(depth < 16, "network config: 'parent' chain is too deep (a cycle?)");
    const std::string parent_path = join_path(dir_of(child_path), child.at("parent").str);
    Json parent = merge_parent_network_config(Json::parse(read_text_file(parent_path, "Parent network config")), parent_path, depth + 1);
    parent.merge_patch(child);
    return parent;
}

void Testbed::reload_network_from_json(const std::string& json, const std::string& config_base_path) {
    m_config = merge_parent_network_config(Json::parse(json), config_base_path);
    m_have_config = true;
    reset_network();
}

static const char* testbed_mode_dir(int mode) { return mode == MODE_TB_NERF ? "nerf" : mode == MODE_TB_SDF ? "sdf" : "image"; }   // to_string(ETestbedMode), common.cu:175-184

// find_network_config (testbed.cu:228-244): the path itself, else <root_dir>/configs/<mode>/<path> for a relative one
std::string Testbed::find_network_config(const std::string& path) const {
    if (path_exists(path) || (!path.empty() && path[0] == '/')) return path;
    const std::string candidate = join_path(join_path(join_path(root_dir, "configs"), testbed_mode_dir(m_mode)), path);   // empty root_dir: ./configs/...
    return path_exists(candidate) ? candidate : path;
}

// reload_network_from_file (testbed.cu:275-309) with load_network_config (:246-273) for a *.json config; snapshots go through load_snapshot
void Testbed::reload_network_from_file(const std::string& path) {
    if (!path.empty()) {                                                         // :276-286: remember the ARGUMENT ("base.json"), not what it resolved to
        const std::string candidate = find_network_config(path);
        if (path_exists(candidate) || !path_exists(network_config_path)) network_config_path = path;
    }
    const std::string full = find_network_config(network_config_path);
    if (path_exists(full)) {
        m_config = merge_parent_network_config(Json::parse(read_text_file(full, "Network config")), full);
    } else if (network_config_path == "base.json") {
        // the reference ships configs/<mode>/base.json beside its binary; a library has no such directory, so the same hyper-parameters are built in
        m_config = Json::parse(builtin_base_config(m_mode));
    } else {
        throw Error("Network config '" + full + "' does not exist.");
    }
    m_have_config = true;
    reset_network();
}

// The hyper-parameters of the reference's default configs (configs/nerf/base.json, configs/sdf/base.json, configs/image/base.json), restricted to
// the keys this library reads: the loss, the Ema / ExponentialDecay / Adam chain, the hash grid and the fully fused MLPs.
const char* Testbed::builtin_base_config(int mode) {
    if (mode == MODE_TB_NERF)
        return R"({"loss":{"otype":"Huber"},
"optimizer":{"otype":"Ema","decay":0.95,"nested":{"otype":"ExponentialDecay","decay_start":20000,"decay_interval":10000,"decay_base":0.33,
 "nested":{"otype":"Adam","learning_rate":1e-2,"beta1":0.9,"beta2":0.99,"epsilon":1e-15,"l2_reg":1e-6}}},
"encoding":{"otype":"HashGrid","n_levels":4,"n_features_per_level":4,"log2_hashmap_size":19,"base_resolution":16},
"network":{"otype":"FullyFusedMLP","activation":"ReLU","output_activation":"None","n_neurons":64,"n_hidden_layers":1},
"dir_encoding":{"otype":"Composite","nested":[{"n_dims_to_encode":3,"otype":"SphericalHarmonics","degree":4},{"otype":"Identity"}]},
"rgb_network":{"otype":"FullyFusedMLP","activation":"ReLU","output_activation":"None","n_neurons":64,"n_hidden_layers":2}})";
    if (mode == MODE_TB_SDF)
        return R"({"loss":{"otype":"MAPE"},
"optimizer":{"otype":"Ema","decay":0.95,"nested":{"otype":"ExponentialDecay","decay_start":10000,"decay_interval":5000,"decay_base":0.33,
 "nested":{"otype":"Adam","learning_rate":1e-4,"beta1":0.9,"beta2":0.99,"epsilon":1e-15,"l2_reg":1e-6}}},
"encoding":{"otype":"HashGrid","n_levels":16,"n_features_per_level":2,"log2_hashmap_size":19,"base_resolution":16},
"network":{"otype":"FullyFusedMLP","activation":"ReLU","output_activation":"None","n_neurons":64,"n_hidden_layers":2}})";
    return R"({"loss":{"otype":"L2"},
"optimizer":{"otype":"ExponentialDecay","decay_start":20000,"decay_interval":10000,"decay_base":0.33,
 "nested":{"otype":"Adam","learning_rate":1e-2,"beta1":0.9,"beta2":0.99,"epsilon":1e-15,"l2_reg":1e-6}},
"encoding":{"otype":"HashGrid","n_levels":16,"n_features_per_level":2,"log2_hashmap_size":24,"base_resolution":16},
"network":{"otype":"FullyFusedMLP","activation":"ReLU","output_activation":"None","n_neurons":64,"n_hidden_layers":2}})";
}

// reset_network (testbed.cu:3903-4151): seeds, counters, network + trainer construction, parameter init
void Testbed::reset_network(bool clear_density_grid) {
    m_rng = Pcg32(seed);                       // testbed.cu:3906
    m_density_grid_rng = Pcg32(m_rng.next_uint());   // testbed.cu:3919 (density_grid_rng seeded from m_rng)
    m_training_step = 0; m_loss_scalar = 0.f;
    m_rays_per_batch = 1u << 12; m_n_rays_total = 0; m_measured = 0; m_measured_before = 0;   // testbed.cu:3912, testbed.h:440
    m_density_grid_ema_step = 0;
    const Json& opt = m_config.contains("optimizer") ? m_config.at("optimizer") : Json::parse("{\"otype\":\"Adam\"}");
    if (m_mode == MODE_TB_NERF) {
        if (m_config.contains("loss")) loss_type = loss_type_from_string(m_config.at("loss").value("otype", "L2"));   // testbed.cu:3953-3956
        m_nerf_net = std::make_unique<NerfNetwork>(m_config);
        m_trainer = std::make_unique<Trainer>(m_nerf_net->n_params, m_nerf_net->n_matrix_params, opt);
        m_nerf_net->initialize_params(seed, m_trainer->master, 0);
        if (m_grid_allocated && clear_density_grid) {
            NGP_CUDA_CHECK(cudaMemset(m_density_grid.ptr, 0, m_density_grid.bytes));
            NGP_CUDA_CHECK(cudaMemset(m_bitfield.ptr, 0, m_bitfield.bytes));
        }
    } else {
        // Trainer(network, optimizer, create_loss(config["loss"])) (testbed.cu:4007,4129); image: 2-D -> rgb, SDF: 3-D -> distance
        m_tcnn_loss = tcnn_loss_from_string(m_config.contains("loss") ? m_config.at("loss").value("otype", "L2") : std::string("L2"));
        m_image_net = m_mode == MODE_TB_SDF ? std::make_unique<EncodedMlp>(m_config, 3, 1) : std::make_unique<EncodedMlp>(m_config, 2, 3);
        m_trainer = std::make_unique<Trainer>(m_image_net->n_params, m_image_net->n_matrix_params, opt);
        m_image_net->initialize_params(seed, m_trainer->master, 0);
    }
    m_trainer->params_from_master(0);
    if (m_mode == MODE_TB_NERF) {   // "Create distortion map model" (testbed.cu:4066-4077): zeros, its own optimizer (the main one's config when none is given)
        m_distortion_res[0] = m_distortion_res[1] = 32;
        const Json* dopt = &opt;
        if (m_config.contains("distortion_map")) {
            const Json& dm = m_config.at("distortion_map");
            if (dm.contains("resolution") && dm.at("resolution").arr.size() == 2) { m_distortion_res[0] = (int)dm.at("resolution").arr[0].num; m_distortion_res[1] = (int)dm.at("resolution").arr[1].num; }
            if (dm.contains("optimizer")) dopt = &dm.at("optimizer");
        }
        NGP_CHECK(m_distortion_res[0] > 0 && m_distortion_res[1] > 0, "distortion_map: bad resolution");
        m_distortion_opt = std::make_unique<Optimizer>(*dopt);
        m_distortion_live = false; ++m_distortion_version;
    }
    NGP_CUDA_CHECK(cudaDeviceSynchronize());
}

void Testbed::ensure_distortion_map(cudaStream_t s) {
    if (m_distortion_live) return;
    const size_t n = (size_t)m_distortion_res[0] * m_distortion_res[1] * 2;
    for (DeviceBuffer* b : {&m_distortion_map, &m_distortion_grad, &m_distortion_weight, &m_distortion_m1, &m_distortion_m2})
        NGP_CUDA_CHECK(cudaMemsetAsync(b->get<float>(n), 0, n * sizeof(float), s));
    NGP_CUDA_CHECK(cudaMemsetAsync(m_distortion_steps.get<uint32_t>(n), 0, n * sizeof(uint32_t), s));
    m_distortion_live = true; ++m_distortion_version;
    m_prefetch_valid = false; m_walk_ahead_valid = false; m_early_prefetch = false;
}

// the distortion part of the camera update (testbed_nerf.cu:3811-3818): gradients / weights, Trainer::optimizer_step(LOSS_SCALE * n_steps_between_cam_updates)
void Testbed::distortion_step(cudaStream_t s) {
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));   // a sampler in flight reads the map
    Optimizer& o = *m_distortion_opt;
    const uint32_t n = (uint32_t)(m_distortion_res[0] * m_distortion_res[1] * 2);
    nerf_distortion_step(s, n, m_distortion_map.as<float>(), m_distortion_grad.as<float>(), m_distortion_weight.as<float>(), m_distortion_m1.as<float>(),
                         m_distortion_m2.as<float>(), m_distortion_steps.as<uint32_t>(), o.current_lr, o.beta1, o.beta2, o.epsilon, LOSS_SCALE * (float)n_steps_between_cam_updates);
    ++o.step_count;
    if (o.decay_interval > 0 && o.step_count >= o.decay_start && o.step_count < o.decay_end && (o.step_count - o.decay_start) % o.decay_interval == 0) o.current_lr *= o.decay_base;
    ++m_distortion_version;
    m_prefetch_valid = false; m_walk_ahead_valid = false; m_early_prefetch = false;
}

// NerfDataset upload + load_nerf_post (testbed_nerf.cu:3046-3112): aabb, cascades, cone angle
void Testbed::set_nerf_dataset(cudaStream_t s, uint32_t n_images, const HostImage* images, int aabb_scale) {
    NGP_CHECK(m_mode == MODE_TB_NERF, "set_nerf_dataset: testbed is not in Nerf mode");
    NGP_CHECK(n_images > 0 && images, "set_nerf_dataset: no images");
    NGP_CHECK(aabb_scale >= 1 && (aabb_scale & (aabb_scale - 1)) == 0, "NeRF dataset's `aabb_scale` must be a power of two");
    NGP_CHECK(aabb_scale <= (1 << (NERF_CASCADES - 1)), "NeRF dataset must have `aabb_scale <= 128`");
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));   // a ray prefetch in flight reads the old images
    if (m_copy_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_copy_stream));
    m_cam_exposure_n = 0; m_n_steps_since_cam_update = 0;   // load_nerf_post (testbed_nerf.cu:3036-3046): fresh exposure optimisers
    m_cdf_valid = false; m_error_map_live = false; m_n_steps_since_error_map_update = 0;   // error map and CDFs belong to the views that were loaded
    m_copy_pending = false;
    m_prefetch_valid = false; m_early_prefetch = false; m_walk_ahead_valid = false; ++m_dataset_version; m_all_images_u8 = true;
    for (void* p : m_image_buffers) cudaFree(p);
    m_image_buffers.clear(); m_image_bytes.clear();
    for (float* p : m_depth_buffers) if (p) cudaFree(p);
    m_depth_buffers.assign(n_images, nullptr);
    m_frame_has_mask.assign(n_images, false);
    std::vector<ImageMeta> metas(n_images);
    for (uint32_t i = 0; i < n_images; ++i) {
        const HostImage& h = images[i];
        NGP_CHECK(h.data_type == IMG_BYTE || h.data_type == IMG_FLOAT, "image data type must be RGBA8 (0) or RGBA float32 (1)");
        NGP_CHECK(h.lens_mode == LENS_PERSPECTIVE || h.lens_mode == LENS_OPENCV, "lens mode must be Perspective (0) or OpenCV (1)");
        const size_t bytes = (size_t)h.res[0] * h.res[1] * (h.data_type == IMG_BYTE ? 4 : 16);
        m_all_images_u8 = m_all_images_u8 && h.data_type == IMG_BYTE;
        if (h.data_type == IMG_BYTE) m_frame_has_mask[i] = has_mask_pixels(h.pixels, (size_t)h.res[0] * h.res[1]);
        void* d = nullptr;
        NGP_CUDA_CHECK(cudaMalloc(&d, bytes));
        m_image_buffers.push_back(d); m_image_bytes.push_back(bytes);
        NGP_CUDA_CHECK(cudaMemcpyAsync(d, h.pixels, bytes, cudaMemcpyHostToDevice, s));
        ImageMeta& m = metas[i];
        m.pixels = d; m.data_type = h.data_type; m.lens_mode = h.lens_mode;
        memcpy(m.lens_params, h.lens_params, sizeof(m.lens_params)); memcpy(m.res, h.res, sizeof(m.res));
        memcpy(m.principal_point, h.principal_point, sizeof(m.principal_point)); memcpy(m.focal_length, h.focal_length, sizeof(m.focal_length));
        memcpy(m.xform, h.xform, sizeof(m.xform)); memcpy(m.xform_end, h.xform, sizeof(m.xform_end));   // start == end, shutter (0,0,0,0)
    }
    // the poses as given; what the kernels see goes through update_transforms (normalisation of scaled rotations, trained offsets)
    m_base_xforms.resize(n_images);
    for (uint32_t i = 0; i < n_images; ++i) { memcpy(m_base_xforms[i].data(), metas[i].xform, 48); memcpy(m_base_xforms[i].data() + 12, metas[i].xform_end, 48); }
    m_cam_pos_offset.assign(n_images, VecAdam{}); m_cam_rot_offset.assign(n_images, VecAdam{});   // load_nerf_post (testbed_nerf.cu:3036-3039)
    m_host_metas = metas; m_n_images = n_images;
    for (uint32_t i = 0; i < n_images; ++i) effective_xform(i, m_host_metas[i].xform, m_host_metas[i].xform_end);
    ImageMeta* md = m_metas.get<ImageMeta>(n_images);
    NGP_CUDA_CHECK(cudaMemcpyAsync(md, m_host_metas.data(), n_images * sizeof(ImageMeta), cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    load_nerf_post(s, aabb_scale);
}

// load_nerf_post (testbed_nerf.cu:3046-3112): aabb, cascades, cone angle; occupancy-grid buffers
void Testbed::load_nerf_post(cudaStream_t s, int aabb_scale) {
    NGP_CHECK(aabb_scale >= 1, "aabb_scale must be at least 1");
    m_aabb_scale = aabb_scale;
    const float half = 0.5f * (float)std::min(1 << (NERF_CASCADES - 1), aabb_scale);   // m_aabb.inflate(0.5 * aabb_scale) around 0.5
    for (int k = 0; k < 3; ++k) { aabb.min[k] = 0.5f - half; aabb.max[k] = 0.5f + half; }
    render_aabb = aabb;
    m_max_cascade = 0;
    while ((1 << m_max_cascade) < aabb_scale) ++m_max_cascade;
    cone_angle_constant = aabb_scale <= 1 ? 0.0f : (1.0f / 256.0f);   // testbed_nerf.cu:3109
    // density grid: always all 8 mips of bitfield (testbed_nerf.cu:3541-3542).  A second scene loaded into a trained testbed (run.py:231 loads the
    // test transforms that way) keeps the network, the training step and - when the cascade count is unchanged, so that GPUMemory::resize is a
    // no-op in the reference (testbed_nerf.cu:3417-3420) - the occupancy grid: load_nerf_post (testbed_nerf.cu:3026-3112) resets none of them
    const size_t n_elements = (size_t)NERF_GRID_N_CELLS * (m_max_cascade + 1);
    const bool keep_grid = m_grid_allocated && m_grid_elements == n_elements;
    if (!keep_grid) {
        NGP_CUDA_CHECK(cudaMemsetAsync(m_density_grid.get<float>(n_elements), 0, n_elements * 4, s));
        m_density_grid_tmp.get<float>(n_elements);
        NGP_CUDA_CHECK(cudaMemsetAsync(m_bitfield.get<uint8_t>((size_t)NERF_GRID_N_CELLS * NERF_CASCADES / 8), 0, (size_t)NERF_GRID_N_CELLS * NERF_CASCADES / 8, s));
        NGP_CUDA_CHECK(cudaMemsetAsync(m_mean.get<float>(4), 0, 16, s));
        m_partials.get<float>(1024);
        m_grid_pos.get<float>(n_elements * 3); m_grid_idx.get<uint32_t>(n_elements); m_grid_out.get<__half>(n_elements);
        m_grid_allocated = true; m_grid_elements = n_elements;
        m_density_grid_ema_step = 0;
    }
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
}

void Testbed::reset(bool reset_density_grid) {
    NGP_CHECK(m_have_config, "reset: no network config loaded");
    NGP_CUDA_CHECK(cudaDeviceSynchronize());
    m_prefetch_valid = false; m_early_prefetch = false; m_walk_ahead_valid = false;
    m_cam_exposure_n = 0; m_n_steps_since_cam_update = 0;   // reset_network -> reset_camera_extrinsics (testbed.cu:3921): exposure optimisers start over
    if (m_n_images && !m_cam_pos_offset.empty()) {          // ... and the pose offsets (testbed_nerf.cu:2921-2933); the given poses are in force again
        m_cam_pos_offset.assign(m_n_images, VecAdam{}); m_cam_rot_offset.assign(m_n_images, VecAdam{});
        update_transforms(0, 0, m_n_images);
    }
    reset_network(reset_density_grid);
}

void Testbed::upload_view_metadata(cudaStream_t s, uint32_t frame_idx) {   // NerfDataset::update_metadata(first, last)
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));   // a ray prefetch in flight reads the old metadata
    m_prefetch_valid = false; m_walk_ahead_valid = false; ++m_dataset_version;
    NGP_CUDA_CHECK(cudaMemcpyAsync(m_metas.as<ImageMeta>() + frame_idx, &m_host_metas[frame_idx], sizeof(ImageMeta), cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
}

void Testbed::set_camera_intrinsics(cudaStream_t s, int frame_idx, float fx, float fy, float cx, float cy, float k1, float k2, float p1, float p2) {
    NGP_CHECK(m_mode == MODE_TB_NERF, "set_camera_intrinsics: testbed is not in Nerf mode");
    if (frame_idx < 0 || frame_idx >= (int)m_n_images) return;   // testbed_nerf.cu:2874-2876
    if (fx <= 0.f) fx = fy;
    if (fy <= 0.f) fy = fx;
    ImageMeta& m = m_host_metas[frame_idx];
    cx = cx < 0.f ? -cx : cx / (float)m.res[0];
    cy = cy < 0.f ? -cy : cy / (float)m.res[1];
    m.lens_mode = LENS_PERSPECTIVE; m.lens_params[0] = m.lens_params[1] = m.lens_params[2] = m.lens_params[3] = 0.f;
    if (k1 != 0.f || k2 != 0.f || p1 != 0.f || p2 != 0.f) { m.lens_mode = LENS_OPENCV; m.lens_params[0] = k1; m.lens_params[1] = k2; m.lens_params[2] = p1; m.lens_params[3] = p2; }
    m.principal_point[0] = cx; m.principal_point[1] = cy; m.focal_length[0] = fx; m.focal_length[1] = fy;
    upload_view_metadata(s, (uint32_t)frame_idx);
}

void Testbed::set_training_depth(cudaStream_t s, int frame_idx, const void* depth_host, int depth_type, float depth_scale) {
    NGP_CHECK(m_mode == MODE_TB_NERF, "set_training_depth: testbed is not in Nerf mode");
    NGP_CHECK(frame_idx >= 0 && frame_idx < (int)m_n_images, "Invalid frame index");
    NGP_CHECK(depth_type == 0 || depth_type == 1, "set_training_depth: depth type must be uint16 (0) or float32 (1)");
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));                                    // no loss kernel is reading the old map
    ImageMeta& m = m_host_metas[frame_idx];
    const size_t n = (size_t)m.res[0] * m.res[1];
    if (depth_scale < 0.f) {                                                     // nerf_loader.cu:957-959: depthmemory[frame_idx].free_memory()
        if (m_depth_buffers[frame_idx]) cudaFree(m_depth_buffers[frame_idx]);
        m_depth_buffers[frame_idx] = nullptr; m.depth = nullptr;
    } else {
        if (!m_depth_buffers[frame_idx]) NGP_CUDA_CHECK(cudaMalloc(&m_depth_buffers[frame_idx], n * sizeof(float)));
        std::vector<float> host(n, 0.f);                                         // copy_depth (nerf_loader.cu:80-90): zeros without data or with scale <= 0
        if (depth_host && depth_scale > 0.f) {
            if (depth_type == 0) { const uint16_t* p = (const uint16_t*)depth_host; for (size_t i = 0; i < n; ++i) host[i] = (float)p[i] * depth_scale; }
            else { const float* p = (const float*)depth_host; for (size_t i = 0; i < n; ++i) host[i] = p[i] * depth_scale; }
        }
        NGP_CUDA_CHECK(cudaMemcpyAsync(m_depth_buffers[frame_idx], host.data(), n * sizeof(float), cudaMemcpyHostToDevice, s));
        NGP_CUDA_CHECK(cudaStreamSynchronize(s));
        m.depth = m_depth_buffers[frame_idx];
    }
    upload_view_metadata(s, (uint32_t)frame_idx);
}

// set_camera_extrinsics_rolling_shutter (testbed_nerf.cu:2896-2915): start and end pose of one view + its rolling-shutter coefficients
void Testbed::set_camera_extrinsics_rolling_shutter(cudaStream_t s, int frame_idx, const float* c2w_start, const float* c2w_end, const float* rolling_shutter4,
                                                    bool convert_to_ngp) {
    NGP_CHECK(m_mode == MODE_TB_NERF && c2w_start && c2w_end && rolling_shutter4, "set_camera_extrinsics_rolling_shutter: testbed is not in Nerf mode / null argument");
    if (frame_idx < 0 || frame_idx >= (int)m_n_images) return;   // :2897-2899
    const float* src[2] = {c2w_start, c2w_end};
    float* dst[2] = {m_base_xforms[frame_idx].data(), m_base_xforms[frame_idx].data() + 12};
    for (int k = 0; k < 2; ++k) {
        float m[12]; memcpy(m, src[k], sizeof(m));
        if (convert_to_ngp) {   // nerf_matrix_to_ngp (nerf_loader.h:120-140)
            for (int r = 0; r < 3; ++r) { m[3 + r] *= -1.f; m[6 + r] *= -1.f; m[9 + r] = m[9 + r] * m_ds_scale + m_ds_offset[r]; }
            float t[12]; memcpy(t, m, sizeof(t));
            for (int c = 0; c < 4; ++c) { m[c * 3 + 0] = t[c * 3 + 1]; m[c * 3 + 1] = t[c * 3 + 2]; m[c * 3 + 2] = t[c * 3 + 0]; }
        }
        memcpy(dst[k], m, sizeof(m));
    }
    memcpy(m_host_metas[frame_idx].rolling_shutter, rolling_shutter4, 16);
    m_cam_pos_offset[frame_idx] = VecAdam{}; m_cam_rot_offset[frame_idx] = VecAdam{};   // :2908-2911 (the exposure of the view is reset there too)
    if (m_cam_exposure_n == m_n_images && m_n_images) {
        NGP_CUDA_CHECK(cudaMemsetAsync(m_cam_exposure.as<float>() + (size_t)frame_idx * 3, 0, 12, s));
        NGP_CUDA_CHECK(cudaMemsetAsync(m_cam_exposure_m1.as<float>() + (size_t)frame_idx * 3, 0, 12, s));
        NGP_CUDA_CHECK(cudaMemsetAsync(m_cam_exposure_m2.as<float>() + (size_t)frame_idx * 3, 0, 12, s));
    }
    update_transforms(s, (uint32_t)frame_idx, (uint32_t)frame_idx + 1);
}

// rotmat / rotvec (common_device.cuh:698-722; glm angle / axis, ext/quaternion_trigonometric.inl:5-29), column-major 3x3
static void rotmat3(const float* v, float* m) {
    const float angle = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (angle == 0.0f) { const float id[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; memcpy(m, id, sizeof(id)); return; }
    const float x = v[0] / angle, y = v[1] / angle, z = v[2] / angle, sn = sinf(angle), c = cosf(angle), oc = 1.0f - c;
    const float r[9] = {oc * x * x + c, oc * x * y + z * sn, oc * z * x - y * sn, oc * x * y - z * sn, oc * y * y + c, oc * y * z + x * sn,
                        oc * z * x + y * sn, oc * y * z - x * sn, oc * z * z + c};
    memcpy(m, r, sizeof(r));
}
static void rotvec3(const float* m, float* v) {
    float q[4]; quat_cast3(m, q);   // (w, x, y, z)
    float angle;
    if (std::fabs(q[0]) > 0.877582561890372716130286068203503191f) {   // cos_one_over_two
        const float a = asinf(std::sqrt(q[1] * q[1] + q[2] * q[2] + q[3] * q[3])) * 2.0f;
        angle = q[0] < 0.0f ? 3.14159265358979323846264338327950288f * 2.0f - a : a;
    } else angle = acosf(q[0]) * 2.0f;
    const float t1 = 1.0f - q[0] * q[0];
    float ax[3] = {0.f, 0.f, 1.f};
    if (t1 > 0.0f) { const float t2 = 1.0f / std::sqrt(t1); ax[0] = q[1] * t2; ax[1] = q[2] * t2; ax[2] = q[3] * t2; }
    for (int k = 0; k < 3; ++k) v[k] = ax[k] * angle;
}
static void mat3_mul(const float* a, const float* b, float* out) {   // column major: out = a * b
    float r[9];
    for (int c = 0; c < 3; ++c) for (int rr = 0; rr < 3; ++rr) r[c * 3 + rr] = a[0 * 3 + rr] * b[c * 3 + 0] + a[1 * 3 + rr] * b[c * 3 + 1] + a[2 * 3 + rr] * b[c * 3 + 2];
    memcpy(out, r, sizeof(r));
}

// NerfDataset xforms -> training transforms (Testbed::Nerf::Training::update_transforms, testbed_nerf.cu:2984-3023): a rotation block with a
// scaling component (|det - 1| > 0.01) is normalised by cbrt(det) (and stays normalised), then the view's trained rotation offset is applied
// from the left and its position offset added
void Testbed::effective_xform(uint32_t i, float* start, float* end) {
    float* base[2] = {m_base_xforms[i].data(), m_base_xforms[i].data() + 12};
    float det[2];
    for (int k = 0; k < 2; ++k) {
        const float* m = base[k];
        det[k] = m[0] * (m[4] * m[8] - m[7] * m[5]) - m[3] * (m[1] * m[8] - m[7] * m[2]) + m[6] * (m[1] * m[5] - m[4] * m[2]);
    }
    if (std::fabs(det[0] - 1.0f) > 0.01f || std::fabs(det[1] - 1.0f) > 0.01f)
        for (int k = 0; k < 2; ++k) { const float c = std::cbrt(det[k]); for (int e = 0; e < 9; ++e) base[k][e] /= c; }
    float rot[9]; rotmat3(m_cam_rot_offset[i].var, rot);
    float* out[2] = {start, end};
    for (int k = 0; k < 2; ++k) {
        mat3_mul(rot, base[k], out[k]);
        for (int r = 0; r < 3; ++r) out[k][9 + r] = base[k][9 + r] + m_cam_pos_offset[i].var[r];
    }
}
void Testbed::update_transforms(cudaStream_t s, uint32_t first, uint32_t last) {
    last = std::min(last, m_n_images);
    if (first >= last) return;
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));   // a ray prefetch in flight reads the old poses
    m_prefetch_valid = false; m_walk_ahead_valid = false; m_early_prefetch = false; ++m_dataset_version;
    for (uint32_t i = first; i < last; ++i) effective_xform(i, m_host_metas[i].xform, m_host_metas[i].xform_end);
    NGP_CUDA_CHECK(cudaMemcpyAsync(m_metas.as<ImageMeta>() + first, &m_host_metas[first], (size_t)(last - first) * sizeof(ImageMeta), cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
}
void Testbed::get_camera_extrinsics(int frame_idx, bool convert_to_ngp, float* out12) const {   // testbed_nerf.cu:2925-2932
    NGP_CHECK(m_mode == MODE_TB_NERF && frame_idx >= 0 && frame_idx < (int)m_host_metas.size(), "Invalid frame index");
    float m[12]; memcpy(m, m_host_metas[frame_idx].xform, sizeof(m));
    if (convert_to_ngp) {   // ngp_matrix_to_nerf (nerf_loader.h:142-158): the inverse of nerf_matrix_to_ngp
        float t[12];
        for (int c = 0; c < 4; ++c) { t[c * 3 + 0] = m[c * 3 + 2]; t[c * 3 + 1] = m[c * 3 + 0]; t[c * 3 + 2] = m[c * 3 + 1]; }
        for (int r = 0; r < 3; ++r) { t[3 + r] *= -1.f; t[6 + r] *= -1.f; t[9 + r] = (t[9 + r] - m_ds_offset[r]) / m_ds_scale; }
        memcpy(m, t, sizeof(m));
    }
    memcpy(out12, m, sizeof(m));
}

// The extrinsics half of the camera update in Testbed::train_nerf (testbed_nerf.cu:3780-3808): gradients to the host, one Adam step per view on
// the position offset (AdamOptimizer<vec3>, adam_optimizer.h:143-150) and on the rotation offset (RotationAdamOptimizer, :238-247: the step is
// composed with the current rotation in matrix form), both with the decaying extrinsic learning rate, then the new poses go to the device
static void vec_adam_moments(Testbed::VecAdam& st, const float* g, float lr, float* delta) {
    ++st.iter;
    const float actual_lr = lr * std::sqrt(1.0f - std::pow(0.99f, (float)st.iter)) / (1.0f - std::pow(0.9f, (float)st.iter));
    for (int k = 0; k < 3; ++k) {
        st.m1[k] = 0.9f * st.m1[k] + (1.0f - 0.9f) * g[k];
        st.m2[k] = 0.99f * st.m2[k] + (1.0f - 0.99f) * g[k] * g[k];
        delta[k] = actual_lr * st.m1[k] / (std::sqrt(st.m2[k]) + 1e-8f);
    }
}
void Testbed::update_error_map_cdfs(cudaStream_t s) {
    m_cdf_res[0] = m_error_map_res[0]; m_cdf_res[1] = m_error_map_res[1];
    const size_t nxy = (size_t)m_cdf_res[0] * m_cdf_res[1] * m_n_images, ny = (size_t)m_cdf_res[1] * m_n_images;
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));        // a sampler in flight may be reading the old CDFs
    float* cxy = m_cdf_x_cond_y.get<float>(nxy); float* cy = m_cdf_y.get<float>(ny); float* cimg = m_cdf_img.get<float>(m_n_images);
    nerf_construct_error_cdfs(s, m_n_images, (uint32_t)m_cdf_res[1], (uint32_t)m_cdf_res[0], m_error_map.as<float>(), cxy, cy, cimg);
    // the view CDF on the host ("single-threaded anyway", :3729-3743): at least MIN_PMF = 0.1 of the mass stays uniform
    std::vector<float> pmf(m_n_images), cdf(m_n_images);
    NGP_CUDA_CHECK(cudaMemcpyAsync(pmf.data(), cimg, m_n_images * sizeof(float), cudaMemcpyDeviceToHost, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    float cum = 0; for (uint32_t i = 0; i < m_n_images; ++i) { cum += pmf[i]; cdf[i] = cum; }
    const float norm = 1.0f / cum, MIN_PMF = 0.1f;
    for (uint32_t i = 0; i < m_n_images; ++i) cdf[i] = (1.0f - MIN_PMF) * cdf[i] * norm + MIN_PMF * (float)(i + 1) / (float)m_n_images;
    NGP_CUDA_CHECK(cudaMemcpyAsync(cimg, cdf.data(), m_n_images * sizeof(float), cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    m_n_steps_since_error_map_update = 0; m_cdf_valid = true; ++m_cdf_version;
    n_steps_between_error_map_updates = (uint32_t)((float)n_steps_between_error_map_updates * 1.5f);
    m_prefetch_valid = false; m_walk_ahead_valid = false; m_early_prefetch = false;   // rays drawn from the old CDFs
}

void Testbed::extrinsics_step(cudaStream_t s) {
    std::vector<float> gp((size_t)m_n_images * 3), gr((size_t)m_n_images * 3);
    NGP_CUDA_CHECK(cudaMemcpyAsync(gp.data(), m_cam_pos_grad.as<float>(), gp.size() * 4, cudaMemcpyDeviceToHost, s));
    NGP_CUDA_CHECK(cudaMemcpyAsync(gr.data(), m_cam_rot_grad.as<float>(), gr.size() * 4, cudaMemcpyDeviceToHost, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    const float per_camera_loss_scale = (float)m_n_images / LOSS_SCALE / (float)n_steps_between_cam_updates;
    const float main_lr = m_trainer->opt.current_lr;
    for (uint32_t i = 0; i < m_n_images; ++i) {
        VecAdam& po = m_cam_pos_offset[i]; VecAdam& ro = m_cam_rot_offset[i];
        float pg[3], rg[3];
        for (int k = 0; k < 3; ++k) {
            pg[k] = gp[(size_t)i * 3 + k] * per_camera_loss_scale + po.var[k] * extrinsic_l2_reg;
            rg[k] = gr[(size_t)i * 3 + k] * per_camera_loss_scale + ro.var[k] * extrinsic_l2_reg;
        }
        const float lr_p = std::max(extrinsic_learning_rate * std::pow(0.33f, (float)(po.iter / 128)), main_lr / 1000.0f);
        const float lr_r = std::max(extrinsic_learning_rate * std::pow(0.33f, (float)(ro.iter / 128)), main_lr / 1000.0f);
        float d[3];
        vec_adam_moments(po, pg, lr_p, d);
        for (int k = 0; k < 3; ++k) po.var[k] -= d[k];
        vec_adam_moments(ro, rg, lr_r, d);
        const float neg[3] = {-d[0], -d[1], -d[2]};
        float a[9], b[9], c[9]; rotmat3(neg, a); rotmat3(ro.var, b); mat3_mul(a, b, c);   // variable = rotvec(rotmat(-rot) * rotmat(variable))
        rotvec3(c, ro.var);
    }
    update_transforms(s, 0, m_n_images);
}



void Testbed::set_camera_extrinsics(cudaStream_t s, int frame_idx, const float* c2w, bool convert_to_ngp) {   // :2917-2919
    const float zero[4] = {0.f, 0.f, 0.f, 0.f};
    NGP_CHECK(m_mode == MODE_TB_NERF && c2w, "set_camera_extrinsics: testbed is not in Nerf mode / null matrix");
    set_camera_extrinsics_rolling_shutter(s, frame_idx, c2w, c2w, zero, convert_to_ngp);
}

void Testbed::set_training_image(cudaStream_t s, uint32_t frame_idx, const void* pixels_host) {
    NGP_CHECK(m_mode == MODE_TB_NERF, "set_training_image: testbed is not in Nerf mode");
    NGP_CHECK(frame_idx < m_n_images, "Invalid frame index");   // python_api.cu:50-52
    NGP_CHECK(pixels_host, "set_training_image: null pixels");
    // A prefetched walk of the next step's rays may be in flight / waiting to be adopted.  It read the OLD pixels, and only to test
    // them against the mask colour: when neither the old nor the new RGBA8 view contains a masked pixel the walk does not depend
    // on the content and stays valid; otherwise (and for float images, where any negative value masks) it is discarded.
    bool content_matters = !m_all_images_u8;
    if (m_all_images_u8) {
        static const bool skip_scan = getenv("NGP_DEBUG_SKIP_MASK_SCAN") != nullptr;   // measurement only: what the host-side scan costs the upload path
        const bool new_mask = skip_scan ? false : has_mask_pixels(pixels_host, m_image_bytes[frame_idx] / 4);
        content_matters = new_mask || m_frame_has_mask[frame_idx];
        m_frame_has_mask[frame_idx] = new_mask;
    }
    if (content_matters) { ++m_dataset_version; m_prefetch_valid = false; m_walk_ahead_valid = false; if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream)); }
    if (!m_all_images_u8) { NGP_CUDA_CHECK(cudaMemcpyAsync(m_image_buffers[frame_idx], pixels_host, m_image_bytes[frame_idx], cudaMemcpyHostToDevice, s)); return; }
    if (!m_copy_stream) {
        NGP_CUDA_CHECK(cudaStreamCreateWithFlags(&m_copy_stream, cudaStreamNonBlocking));
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_copy_gate, cudaEventDisableTiming));
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_copy_done, cudaEventDisableTiming));
    }
    NGP_CUDA_CHECK(cudaEventRecord(m_ev_copy_gate, s));                          // readers of the old pixels already queued on s
    NGP_CUDA_CHECK(cudaStreamWaitEvent(m_copy_stream, m_ev_copy_gate, 0));
    NGP_CUDA_CHECK(cudaMemcpyAsync(m_image_buffers[frame_idx], pixels_host, m_image_bytes[frame_idx], cudaMemcpyHostToDevice, m_copy_stream));
    NGP_CUDA_CHECK(cudaEventRecord(m_ev_copy_done, m_copy_stream));
    m_copy_pending = true;                                                       // the next loss kernel (the reader of pixel values) waits for it
}

void Testbed::set_image(cudaStream_t s, const float* rgba_host, int w, int h) {
    NGP_CHECK(m_mode == MODE_TB_IMAGE, "set_image: testbed is not in Image mode");
    NGP_CHECK(rgba_host && w > 1 && h > 1, "set_image: bad image");
    NGP_CUDA_CHECK(cudaMemcpyAsync(m_image.get<float>((size_t)w * h * 4), rgba_host, (size_t)w * h * 16, cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    m_image_w = w; m_image_h = h;
}

// load_stbi_image for an 8-bit file (testbed_image.cu:404-418, load_stbi_gpu common_device.cu:47-76): the decoded bytes go up as they are and
// become linear premultiplied floats on the device
void Testbed::set_image_rgba8(cudaStream_t s, const uint8_t* rgba8_host, int w, int h) {
    NGP_CHECK(m_mode == MODE_TB_IMAGE, "set_image: testbed is not in Image mode");
    NGP_CHECK(rgba8_host && w > 1 && h > 1, "set_image: bad image");
    const size_t n = (size_t)w * h;
    DeviceBuffer bytes;
    NGP_CUDA_CHECK(cudaMemcpyAsync(bytes.get<uint8_t>(n * 4), rgba8_host, n * 4, cudaMemcpyHostToDevice, s));
    image_from_rgba8(s, n, bytes.as<uint8_t>(), m_image.get<float>(n * 4));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    m_image_w = w; m_image_h = h;
}

// ---------------------------------------------------------------------------------------------------------------
// update_density_grid_nerf (testbed_nerf.cu:3412-3536) + training_prep_nerf (:4137-4152)
// ---------------------------------------------------------------------------------------------------------------
void Testbed::update_density_grid_nerf(cudaStream_t s, float decay, uint32_t n_uniform, uint32_t n_nonuniform) {
    NGP_PROFILE_SCOPE(s, "density_grid_update");
    const uint32_t n_cascades = m_max_cascade + 1;
    const uint32_t n_elements = NERF_GRID_N_CELLS * n_cascades;
    const uint32_t n_samples = n_uniform + n_nonuniform;
    float* grid = m_density_grid.as<float>(); float* tmp = m_density_grid_tmp.as<float>();
    float* pos = m_grid_pos.as<float>(); uint32_t* idx = m_grid_idx.as<uint32_t>(); __half* out = m_grid_out.as<__half>();
    if (m_training_step == 0) {
        m_density_grid_ema_step = 0;
        nerf_mark_untrained_grid(s, n_elements, grid, m_n_images, m_metas.as<ImageMeta>(), true);
    }
    NGP_CUDA_CHECK(cudaMemsetAsync(tmp, 0, sizeof(float) * n_elements, s));
    nerf_generate_grid_samples(s, n_uniform, m_density_grid_rng, m_density_grid_ema_step, aabb, grid, pos, idx, n_cascades, -0.01f);
    m_density_grid_rng.advance();
    nerf_generate_grid_samples(s, n_nonuniform, m_density_grid_rng, m_density_grid_ema_step, aabb, grid, pos + (size_t)n_uniform * 3, idx + n_uniform, n_cascades,
                               NERF_MIN_OPTICAL_THICKNESS);
    m_density_grid_rng.advance();
    // density() on the TRAINING params (use_inference_params = false, :3514), chunks of 2*128^3
    for (uint32_t i = 0; i < n_samples; i += NERF_GRID_N_CELLS * 2) {
        const uint32_t batch = std::min(NERF_GRID_N_CELLS * 2, n_samples - i);
        m_nerf_net->density(s, m_trainer->params, pos + (size_t)i * 3, 3, batch, out + i);
    }
    nerf_splat_grid_samples(s, n_samples, idx, out, tmp, density_activation);
    nerf_ema_grid(s, n_elements, decay, grid, tmp);
    ++m_density_grid_ema_step;
    nerf_update_bitfield(s, grid, n_cascades, m_bitfield.as<uint8_t>(), m_mean.as<float>(), m_partials.as<float>());
    nerf_build_coarse_occupancy(s, m_bitfield.as<uint8_t>(), m_coarse.get<uint32_t>(1024));
}

void Testbed::training_prep_nerf(cudaStream_t s) {
    const uint32_t n_cascades = m_max_cascade + 1;
    if (m_training_step < 256) update_density_grid_nerf(s, density_grid_decay, NERF_GRID_N_CELLS * n_cascades, 0);
    else update_density_grid_nerf(s, density_grid_decay, NERF_GRID_N_CELLS / 4 * n_cascades, NERF_GRID_N_CELLS / 4 * n_cascades);
}

// ---------------------------------------------------------------------------------------------------------------
// Testbed::train (testbed.cu:4285-4370)
// ---------------------------------------------------------------------------------------------------------------
void Testbed::train(cudaStream_t s, uint32_t batch_size) { train_begin(s, batch_size); train_end(s, 0, 0); }

void Testbed::train_begin(cudaStream_t s, uint32_t batch_size) {
    finish_pending_step();
    if (!m_trainer) reload_network_from_file();          // testbed.cu:4297-4306: training data without an explicitly loaded network -> the default config
    NGP_CHECK(m_trainer, "Unable to create a neural network trainer.");
    NGP_CHECK(batch_size >= BATCH_GRANULARITY && batch_size % BATCH_GRANULARITY == 0, "training batch size must be a multiple of 256");
    if (m_mode == MODE_TB_NERF) {
        NGP_CHECK(m_n_images > 0, "Cannot train without training data");
        if (optimize_distortion) ensure_distortion_map(s);   // before this step's sampler looks at it (a prefetched walk without the map is discarded)
        const uint32_t n_prep_to_skip = std::min(std::max(m_training_step / 16u, 1u), 16u);
        if (m_training_step % n_prep_to_skip == 0) training_prep_nerf(s);
    }
    // optimizer->update_hyperparams (testbed.cu:4339-4346)
    m_trainer->opt.optimize_matrix = shall_train_network; m_trainer->opt.optimize_non_matrix = shall_train_encoding;
    m_get_loss = m_training_step % 16 == 0;
    m_pending_batch = batch_size;
    if (m_mode == MODE_TB_NERF) train_nerf_step(s, batch_size);
    else if (m_mode == MODE_TB_SDF) train_sdf_step(s, batch_size, m_get_loss);
    else train_image_step(s, batch_size, m_get_loss);
}

void Testbed::comm_init(const void* id128, uint32_t rank, uint32_t world, const char* nccl_lib_path) {
    NGP_CHECK(m_mode == MODE_TB_NERF, "comm_init: data-parallel training is implemented for the Nerf mode");
    NGP_CHECK(world >= 1 && rank < world, "comm_init: bad rank / world size");
    m_comm = std::make_unique<Comm>(id128, (int)rank, (int)world, nccl_lib_path);
    set_shard(rank, world);
    if (!m_comm_stream) {
        int least = 0, greatest = 0;
        NGP_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        NGP_CUDA_CHECK(cudaStreamCreateWithPriority(&m_comm_stream, cudaStreamNonBlocking, greatest));   // the collective's CTAs get their SMs first
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_slice, cudaEventDisableTiming));
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_comm_done, cudaEventDisableTiming));
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_dp_counters, cudaEventDisableTiming));
    }
    m_dp_counters.get<uint32_t>(2);
    m_prefetch_valid = false;
}

bool Testbed::sharded_step() const {
    return dp_sharded_optimizer && m_comm && m_trainer && m_comm->world() > 1 && m_trainer->n_cap % (uint64_t)m_comm->world() == 0 &&
           (m_trainer->n_cap / (uint64_t)m_comm->world()) % 128 == 0;
}

void Testbed::sync_inference_params(cudaStream_t s) {
    if (!m_comm || !m_trainer || !m_inference_stale) return;
    if (m_trainer->inference_params != m_trainer->params)
        m_comm->all_gather(m_trainer->inference_params, m_trainer->n_cap / (uint64_t)m_comm->world(), COMM_F16, s);
    m_inference_stale = false;
}

void Testbed::pending_counters(uint32_t* measured, uint32_t* measured_before) {
    NGP_CHECK(m_pending_batch > 0 && m_mode == MODE_TB_NERF, "pending_counters: no Nerf training step in flight (call train_begin first)");
    NGP_CUDA_CHECK(cudaEventSynchronize(m_ev_counters));
    *measured = m_host_counters[2]; *measured_before = m_host_counters[1];
}

void Testbed::train_end(cudaStream_t s, uint32_t measured_override, uint32_t measured_before_override) {
    NGP_CHECK(m_pending_batch > 0, "train_end without train_begin");
    const uint32_t target = m_pending_batch; m_pending_batch = 0;
    if (m_mode == MODE_TB_SDF && m_sdf_step_skipped) return;   // train_sdf does nothing when the data set is smaller than the batch (testbed_sdf.cu:1293)
    if (m_mode == MODE_TB_NERF && sharded_step()) {
        // reduce-scatter happened in train_nerf_step: this rank's slice of `grads` holds the global sums
        const uint64_t shard = m_trainer->n_cap / (uint64_t)m_comm->world();
        m_trainer->optimizer_step_shard(s, LOSS_SCALE, shard * m_comm->rank(), shard);
        m_comm->all_gather(m_trainer->params, shard, COMM_F16, s);      // the next forward pass (ordered behind it on s) reads every slice
        m_inference_stale = true;
    } else {
        m_trainer->optimizer_step(s, LOSS_SCALE);   // testbed_nerf.cu:3678 / testbed_image.cu:283
    }
    ++m_training_step;
    if (m_mode == MODE_TB_IMAGE || m_mode == MODE_TB_SDF) {
        // the host waits only for the loss read-back (every 16th step, testbed.cu:4349-4353); otherwise it runs ahead of the device
        if (m_get_loss) { NGP_CUDA_CHECK(cudaMemcpyAsync(&m_loss_scalar, m_loss_sum.as<float>(), 4, cudaMemcpyDeviceToHost, s)); NGP_CUDA_CHECK(cudaStreamSynchronize(s)); }
        return;
    }
    // the host half may wait: the next sampler is already in flight behind the device-side law, and nothing else here touches the device
    static const bool defer_env = [] { const char* e = getenv("NGP_DEFER_FINISH"); return e ? atoi(e) != 0 : true; }();
    if (deferred_step_finish && defer_env && m_early_prefetch && !m_comm && !measured_override && !measured_before_override && !walk_ahead &&
        !optimize_exposure && !m_error_map_live) {
        m_finish_pending = true; m_finish_target = target; m_finish_stream = s;
        return;
    }
    finish_step(s, target, measured_override, measured_before_override);
}

void Testbed::finish_pending_step() {
    if (!m_finish_pending) return;
    m_finish_pending = false;
    finish_step(m_finish_stream, m_finish_target, 0, 0);
}

void Testbed::finish_step(cudaStream_t s, uint32_t target, uint32_t measured_override, uint32_t measured_before_override) {
    // NerfCounters::update_after_training (testbed_nerf.cu:3583-3609): two counters (+ the loss every 16 steps) read back.
    // The counters were copied to pinned memory right after the loss kernel (train_nerf_step), so the host has them while
    // the backward pass and the optimizer are still running: time enough to start the next step's rays beside them.
    // (the loss sum of every 16th step was reduced and copied right behind the loss kernel, train_nerf_step: it arrives with the counters)
    NGP_CUDA_CHECK(cudaEventSynchronize(m_ev_counters));
    const uint32_t* c = m_host_counters;
    if (m_comm && !measured_override && !measured_before_override) {   // lock-step control law: the mean counts over the ranks
        NGP_CUDA_CHECK(cudaEventSynchronize(m_ev_dp_counters));
        measured_before_override = std::max(c[5] / (uint32_t)m_comm->world(), 1u);
        measured_override = std::max(c[6] / (uint32_t)m_comm->world(), 1u);
    }
    m_measured_before = measured_before_override ? measured_before_override : c[1];
    m_measured = measured_override ? measured_override : c[2];
    if (m_measured_before == 0 || m_measured == 0) {
        m_measured = m_measured_before = 0; m_loss_scalar = 0.f;
        cudaStreamSynchronize(s);
        throw Error("Nerf training generated 0 samples. Aborting training.");   // testbed_nerf.cu:3693-3697
    }
    m_measured_total += m_measured; m_measured_before_total += m_measured_before;
    m_rays_per_batch = (uint32_t)((float)m_rays_per_batch * (float)target / (float)m_measured);
    m_rays_per_batch = std::min((uint32_t)next_multiple(m_rays_per_batch, BATCH_GRANULARITY), 1u << 18);
    // camera-side optimisers (testbed_nerf.cu:3754-3861; of those this library trains the exposure)
    // "Compute CDFs from the error map" (testbed_nerf.cu:3700-3751)
    if (m_error_map_live) {
        m_n_steps_since_error_map_update += 1;
        if (m_n_steps_since_error_map_update >= n_steps_between_error_map_updates) update_error_map_cdfs(s);
    }
    m_n_steps_since_cam_update += 1;
    const bool train_exposure = optimize_exposure && m_cam_exposure_n == m_n_images, train_extrinsics = optimize_extrinsics && m_cam_pos_grad.ptr != nullptr;
    const bool train_distortion = optimize_distortion && m_distortion_live;
    if ((train_exposure || train_extrinsics || train_distortion) && m_n_steps_since_cam_update >= n_steps_between_cam_updates) {   // :3767-3771
        if (train_extrinsics) extrinsics_step(s);   // host-side Adam per view, as in the reference; the new poses are on the device before the next sampler runs
        if (train_distortion) distortion_step(s);
        if (train_exposure) {
            const float per_camera_loss_scale = (float)m_n_images / LOSS_SCALE / (float)n_steps_between_cam_updates;
            nerf_exposure_step(s, m_n_images, m_cam_exposure_grad.as<float>(), m_cam_exposure.as<float>(), m_cam_exposure_m1.as<float>(), m_cam_exposure_m2.as<float>(),
                               ++m_cam_exposure_iter, m_trainer->opt.current_lr, exposure_l2_reg, per_camera_loss_scale);
        }
        m_n_steps_since_cam_update = 0;
    }
    if (m_walk_ahead_valid) {
        // the counting walk of the next step ran (or is running) with a predicted ray count: right -> only the cheap second phase is left
        if (!adopt_walk_ahead(target)) launch_ray_prefetch(s, target);
    } else if (m_early_prefetch) {
        // the next step's sampler is already running (launch_early_prefetch): it is valid if the law evaluated here gives counts within the
        // bounds its grids were sized for (the device evaluated the same arithmetic and generated nothing otherwise)
        m_early_prefetch = false;
        const uint32_t mi = max_inference_samples(target);
        m_prefetch_key = sampler_key(m_rays_per_batch, mi, target);
        m_prefetch_valid = m_rays_per_batch <= m_early_rays_bound && mi <= m_early_samples_bound && m_early_dataset_version == m_dataset_version;
    } else {
        launch_ray_prefetch(s, target);
    }
    launch_walk_ahead(target);
    // No host wait for the optimizer: the control law only needed the counters (already here), and everything that consumes
    // the new weights is ordered behind them on `s`.  The host so runs half a step ahead and the device never idles between
    // steps.  The loss of every 16th step came with the counters.
    if (m_get_loss) {
        float loss_sum; memcpy(&loss_sum, m_host_counters + 4, 4); m_loss_scalar = loss_sum * (float)m_measured / (float)target;
        // data parallel: every ray's loss is divided by the GLOBAL ray count but only this rank's rays are summed here, so the sum is
        // ~1/world of the step's loss; scaled back it is this rank's estimate of the global mean (comparable with single-GPU runs)
        m_loss_scalar *= (float)m_world;
        memcpy(&m_density_grid_mean, m_host_counters + 7, 4);
    }
}

TrainStats Testbed::stats() const {
    TrainStats st{};
    st.training_step = m_training_step; st.loss = m_loss_scalar; st.rays_per_batch = m_rays_per_batch; st.measured_batch_size = m_measured;
    st.measured_batch_size_before_compaction = m_measured_before; st.n_rays_total = m_n_rays_total;
    st.learning_rate = m_trainer ? m_trainer->opt.current_lr : 0.f;
    st.density_grid_mean = m_density_grid_mean;
    return st;
}


// ---------------------------------------------------------------------------------------------------------------
// train_nerf_step (testbed_nerf.cu:3867-4132)
// ---------------------------------------------------------------------------------------------------------------
Testbed::SamplerKey Testbed::sampler_key(uint32_t n_rays, uint32_t max_inference, uint32_t target) const {
    return {m_training_step, n_rays, max_inference, target, m_rank, m_world, sampler_lanes_per_ray, m_dataset_version, m_rng.state, m_rng.inc,
            cone_angle_constant, train_snap_to_pixel_centers, max_level_rand_training,
            // the CDFs the sampler draws from: their version, and which of the two switches use them
            m_cdf_valid ? (m_cdf_version * 4u + (sample_focal_plane_proportional_to_error ? 1u : 0u) + (sample_image_proportional_to_error ? 2u : 0u)) : 0u,
            m_distortion_live ? m_distortion_version + 1u : 0u};
}

SamplerBuffers Testbed::sampler_buffers(cudaStream_t s, uint32_t set, uint32_t target, uint32_t n_rays) {
    const uint32_t max_samples = target * 16;   // "somewhat of a worst case" (:3873)
    const uint32_t max_rays = 1u << 18;
    const uint32_t max_blocks = max_rays / 8 + 1;   // 256 / lanes_per_ray rays per block, lanes_per_ray <= 32
    SamplerBuffers sb{};
    sb.max_level_rand = max_level_rand_training ? 1u : 0u;
    sb.distortion = distortion_view();
    if (m_cdf_valid) {   // :3963-3967,4027-4031: each switch hands its CDFs to the sampler and to every kernel that replays the ray's rng stream
        if (sample_focal_plane_proportional_to_error) { sb.cdf.cdf_x_cond_y = m_cdf_x_cond_y.as<float>(); sb.cdf.cdf_y = m_cdf_y.as<float>(); }
        if (sample_image_proportional_to_error) sb.cdf.cdf_img = m_cdf_img.as<float>();
        sb.cdf.res[0] = m_cdf_res[0]; sb.cdf.res[1] = m_cdf_res[1];
    }
    sb.numsteps_all = m_numsteps_all[set].get<uint32_t>(max_rays); sb.block_steps = m_block_steps[set].get<uint32_t>(max_blocks);
    sb.block_rays = m_block_rays[set].get<uint32_t>(max_blocks); sb.ray_tmp = m_ray_tmp[set].get<float>((size_t)max_rays * 8);
    sb.ray_indices = m_ray_indices[set].get<uint32_t>(max_rays); sb.rays = m_rays[set].get<float>((size_t)max_rays * 6);
    sb.numsteps = m_numsteps[set].get<uint32_t>((size_t)max_rays * 2);
    const bool fresh_coords = m_coords[set].bytes < (size_t)max_samples * 7 * 4;
    sb.coords = m_coords[set].get<float>((size_t)max_samples * 7);
    if (fresh_coords) NGP_CUDA_CHECK(cudaMemsetAsync(sb.coords, 0, (size_t)max_samples * 7 * 4, s));   // stale rows are fed to the network: keep them finite
    sb.counters = m_counters[set].get<uint32_t>(8);
    sb.cand_list = m_cand_list[set].get<uint32_t>(max_rays);
    // classification needs the unit-cube scene (one cascade); otherwise every ray is walked
    const bool unit = m_max_cascade == 0 && aabb.min[0] == 0.f && aabb.min[1] == 0.f && aabb.min[2] == 0.f && aabb.max[0] == 1.f && aabb.max[1] == 1.f && aabb.max[2] == 1.f;
    sb.coarse = unit && m_coarse.ptr ? m_coarse.as<uint32_t>() : nullptr;
    sb.lanes_per_ray = sampler_lanes_per_ray;
    sb.sample_scratch = m_sample_scratch[set].get<float2>(max_samples);
    sb.scratch_per_ray = std::min<uint32_t>(NERF_STEPS, max_samples / n_rays);
    sb.pool_slots = 4096; sb.sample_pool = m_sample_pool[set].get<float2>((size_t)sb.pool_slots * NERF_STEPS);
    sb.overflow_slot = m_overflow_slot[set].get<uint32_t>(max_rays);
    return sb;
}

uint32_t Testbed::max_inference_samples(uint32_t target) {
    const uint32_t max_samples = target * 16;
    if (m_measured_before == 0) return m_measured_before = max_samples;
    return (uint32_t)next_multiple(std::min(m_measured_before, max_samples), BATCH_GRANULARITY);
}

// prepare_for_training_steps (:3569-3581) + generate_training_samples_nerf (:1382-1658)
void Testbed::launch_sampler(cudaStream_t s, const SamplerKey& key, const SamplerBuffers& sb) {
    NGP_CUDA_CHECK(cudaMemsetAsync(sb.counters, 0, 32, s));
    nerf_generate_training_samples(s, key.n_rays, aabb, key.max_inference, 0, m_rng, m_n_images, m_metas.as<ImageMeta>(), m_bitfield.as<uint8_t>(), m_max_cascade,
                                   train_snap_to_pixel_centers, cone_angle_constant, m_rank * key.n_rays, m_world * key.n_rays, sb);
}

// The rays of step N + 1 depend on the rng, the control law's rays_per_batch (known once step N's loss has run) and the
// occupancy bitfield — not on the weights.  When step N + 1 will not refresh the bitfield first, its walk is started here
// on a second stream, beside step N's backward pass / optimizer (latency-bound traversal next to HBM- and atomics-bound
// kernels), into the buffer set step N does not use.  train_nerf_step(N + 1) adopts it only if every input still matches.
void Testbed::launch_ray_prefetch(cudaStream_t s, uint32_t target) {
    if (!prefetch_rays || m_mode != MODE_TB_NERF || !m_all_images_u8) return;
    const uint32_t n_prep_to_skip = std::min(std::max(m_training_step / 16u, 1u), 16u);
    if (m_training_step % n_prep_to_skip == 0) return;      // the next step starts with training_prep_nerf (new bitfield)
    const uint32_t R = m_rays_per_batch, set = m_training_step % N_SETS;
    const SamplerKey key = sampler_key(R, max_inference_samples(target), target);
    // start beside the hash-grid backward pass and the optimizer (atomics- and HBM-bound, few issue slots used); the fused MLP
    // backward kernel (latency-bound, 8 warps per SM) loses more from sharing its SMs than the overlap wins
    // That holds for the flat marcher (aabb_scale 1: ~65 M warp instructions at 60 % issue utilisation).  The general marcher
    // (cascades, cone angle) is one long dependent chain per ray at ~2 warps per scheduler: it takes longer than the whole rest
    // of the step and uses a tenth of the issue slots, so it starts as soon as the counters are known.
    const bool flat = cone_angle_constant <= 1e-5f && m_max_cascade == 0;
    if (flat) NGP_CUDA_CHECK(cudaStreamWaitEvent(m_aux_stream, m_ev_after_mlp, 0));
    const SamplerBuffers sb = sampler_buffers(m_aux_stream, set, target, R);
    launch_sampler(m_aux_stream, key, sb);
    NGP_CUDA_CHECK(cudaEventRecord(m_ev_prefetch, m_aux_stream));
    m_prefetch_key = key; m_prefetch_valid = true; m_prefetch_inflight = true;
}

void Testbed::ensure_cam_exposure(cudaStream_t s) {
    if (m_cam_exposure_n == m_n_images) return;
    for (DeviceBuffer* b : {&m_cam_exposure, &m_cam_exposure_grad, &m_cam_exposure_m1, &m_cam_exposure_m2})
        NGP_CUDA_CHECK(cudaMemsetAsync(b->get<float>((size_t)m_n_images * 3), 0, (size_t)m_n_images * 12, s));
    m_cam_exposure_iter = 0; m_cam_exposure_n = m_n_images;
}

// Walk-ahead.  With cascades and a cone angle the counting walk is one long dependent chain per ray (fox: 0.8 ms, more than the whole rest of
// the step) and sits on the cycle loss(N) -> rays_per_batch(N+1) -> walk(N+1) -> forward(N+1) -> loss(N+1): prefetching one step ahead hides
// it behind the backward pass only.  What the walk of step N+2 needs — the bitfield, the rng, the images and the RAY COUNT — is all known a
// step early except the ray count, and that is a multiple of 256 which the control law moves rarely once training has settled.  So, called
// from train_end(N) (m_training_step == N+1 there), this enqueues the walk phase of step N+2 on the auxiliary stream behind the prefetch of
// step N+1, predicting rays_per_batch(N+2) == rays_per_batch(N+1).  train_end(N+1) knows the true count: equal -> adopt_walk_ahead launches
// the cheap emit phase with the true sample cap; different (or any other input changed) -> the walk is discarded and the prefetch runs
// as before.  A third buffer set keeps step N's outputs readable meanwhile.
void Testbed::launch_walk_ahead(uint32_t target) {
    if (!walk_ahead || !prefetch_rays || m_mode != MODE_TB_NERF || !m_all_images_u8 || m_comm || m_world != 1) return;
    if (cone_angle_constant <= 1e-5f && m_max_cascade == 0) return;     // flat marcher: the walk is short and already hidden
    if (!m_prefetch_valid) return;                                        // nothing in flight for step N+1: no steady state yet
    const uint32_t s1 = m_training_step, s2 = m_training_step + 1u;       // N+1, N+2
    const auto refreshes = [](uint32_t step) { const uint32_t skip = std::min(std::max(step / 16u, 1u), 16u); return step % skip == 0; };
    if (refreshes(s1) || refreshes(s2)) return;                           // a new bitfield would appear under, or after, the walk
    const uint32_t R = m_rays_per_batch;
    const uint32_t lanes = sampler_lanes_per_ray ? sampler_lanes_per_ray : nerf_sampler_lanes_per_ray(R);
    if (lanes == 1) return;
    if (!m_ev_walk_ahead) NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_walk_ahead, cudaEventDisableTiming));
    Pcg32 rng = m_rng; rng.advance();                                     // the rng step N+2 will sample with (advanced once per step, :4127)
    SamplerKey key = sampler_key(R, 0, target);
    key.step = s2; key.rng_state = rng.state; key.rng_inc = rng.inc;
    const SamplerBuffers sb = sampler_buffers(m_aux_stream, s2 % N_SETS, target, R);
    NGP_CUDA_CHECK(cudaMemsetAsync(sb.counters, 0, 32, m_aux_stream));
    nerf_generate_training_samples(m_aux_stream, R, aabb, 0, 0, rng, m_n_images, m_metas.as<ImageMeta>(), m_bitfield.as<uint8_t>(), m_max_cascade,
                                   train_snap_to_pixel_centers, cone_angle_constant, 0, R, sb, SAMPLER_WALK);
    NGP_CUDA_CHECK(cudaEventRecord(m_ev_walk_ahead, m_aux_stream));
    m_walk_ahead_key = key; m_walk_ahead_rng = rng; m_walk_ahead_valid = true; m_walk_ahead_inflight = true;
}

// train_end(N+1): m_training_step == N+2, m_rays_per_batch and the sample cap of step N+2 are known
bool Testbed::adopt_walk_ahead(uint32_t target) {
    m_walk_ahead_valid = false;
    const uint32_t R = m_rays_per_batch, mi = max_inference_samples(target);
    SamplerKey want = sampler_key(R, 0, target);
    if (!(want == m_walk_ahead_key) || !prefetch_rays) { ++walk_ahead_misses; return false; }
    const SamplerKey key = sampler_key(R, mi, target);
    const SamplerBuffers sb = sampler_buffers(m_aux_stream, m_training_step % N_SETS, target, R);
    nerf_generate_training_samples(m_aux_stream, R, aabb, mi, 0, m_rng, m_n_images, m_metas.as<ImageMeta>(), m_bitfield.as<uint8_t>(), m_max_cascade,
                                   train_snap_to_pixel_centers, cone_angle_constant, 0, R, sb, SAMPLER_EMIT);
    NGP_CUDA_CHECK(cudaEventRecord(m_ev_prefetch, m_aux_stream));
    m_prefetch_key = key; m_prefetch_valid = true; m_prefetch_inflight = true; m_walk_ahead_inflight = false;   // the emit phase is ordered behind the walk
    ++walk_ahead_hits;
    return true;
}

// The same prefetch, enqueued from train_nerf_step(N) BEFORE the host has seen step N's counters: nerf_control_law (already on `s`, right
// behind the loss kernel) leaves rays_per_batch / the inference size of step N+1 in device memory and the sampler kernels read them from
// there; their grids are sized for the bounds.  train_end evaluates the same law on the host and adopts the result if it fits the bounds.
void Testbed::launch_early_prefetch(cudaStream_t s, uint32_t target, uint32_t rays_bound, uint32_t samples_bound) {
    (void)s;
    const uint32_t set = (m_training_step + 1u) % N_SETS;      // m_training_step is still this step's; m_rng has already been advanced
    static const bool after_mlp = [] { const char* e = getenv("NGP_PREFETCH_AFTER"); return e && std::string(e) == "mlp"; }();
    NGP_CUDA_CHECK(cudaStreamWaitEvent(m_aux_stream, after_mlp ? m_ev_after_mlp : m_ev_after_loss, 0));
    SamplerBuffers sb = sampler_buffers(m_aux_stream, set, target, rays_bound);
    sb.dyn = m_step_params[set].as<StepParams>();
    NGP_CUDA_CHECK(cudaMemsetAsync(sb.counters, 0, 32, m_aux_stream));
    nerf_generate_training_samples(m_aux_stream, rays_bound, aabb, samples_bound, 0, m_rng, m_n_images, m_metas.as<ImageMeta>(), m_bitfield.as<uint8_t>(), m_max_cascade,
                                   train_snap_to_pixel_centers, cone_angle_constant, 0, rays_bound, sb);
    NGP_CUDA_CHECK(cudaEventRecord(m_ev_prefetch, m_aux_stream));
    m_early_prefetch = true; m_early_rays_bound = rays_bound; m_early_samples_bound = samples_bound; m_prefetch_inflight = true;
    m_early_dataset_version = m_dataset_version;
}

void Testbed::train_nerf_step(cudaStream_t s, uint32_t target) {
    NGP_PROFILE_SCOPE(s, "train_nerf_step");
    if (!m_aux_stream) {
        int least = 0, greatest = 0;
        NGP_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        NGP_CUDA_CHECK(cudaStreamCreateWithPriority(&m_aux_stream, cudaStreamNonBlocking, least));   // never ahead of the training stream
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_counters, cudaEventDisableTiming));
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_prefetch, cudaEventDisableTiming));
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_after_mlp, cudaEventDisableTiming));
        NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_after_loss, cudaEventDisableTiming));
        NGP_CUDA_CHECK(cudaMallocHost(&m_host_counters, 8 * sizeof(uint32_t)));
        memset(m_host_counters, 0, 8 * sizeof(uint32_t));
    }
    const uint32_t max_samples = target * 16;
    const uint32_t R = m_rays_per_batch, max_rays = 1u << 18;
    const uint32_t set = m_training_step % N_SETS;
    m_last_set = set;
    const SamplerBuffers sb = sampler_buffers(s, set, target, R);
    LossBuffers lb{};
    __half* mlp_out = m_mlp_out.get<__half>((size_t)max_samples * 4);
    lb.network_output = mlp_out; lb.dloss_doutput = m_dloss.get<__half>((size_t)target * 4);
    lb.coords_compacted = m_coords_compacted.get<float>((size_t)target * 7); lb.loss = m_loss.get<float>(max_rays);
    lb.ray_scratch = m_ray_scratch.get<float>((size_t)max_rays * RAY_SCRATCH); lb.compacted_numsteps = m_compacted_numsteps.get<uint32_t>(max_rays);
    m_loss_sum.get<float>(4);
    lb.t_before = m_t_before.get<float>(max_samples);
    lb.enc_width = m_nerf_net->dims.pos_width; lb.enc_out = m_nerf_net->training_encoding(target);

    const uint32_t max_inference = max_inference_samples(target);
    if (m_training_step == 0) m_n_rays_total = 0;
    const uint32_t n_rays_total = m_n_rays_total;
    m_n_rays_total += R;

    const SamplerKey key = sampler_key(R, max_inference, target);
    const bool prefetched = m_prefetch_valid && m_prefetch_key == key;
    if (m_prefetch_valid || m_prefetch_inflight) NGP_CUDA_CHECK(cudaStreamWaitEvent(s, m_ev_prefetch, 0));   // adopted or stale, it must have left the buffers
    m_prefetch_valid = false; m_prefetch_inflight = false;
    if (prefetched) ++prefetch_hits;
    else {
        if (m_walk_ahead_inflight) { NGP_CUDA_CHECK(cudaStreamWaitEvent(s, m_ev_walk_ahead, 0)); m_walk_ahead_inflight = false; }   // a discarded walk may still be writing one of the sets
        launch_sampler(s, key, sb);
    }
    NGP_CUDA_CHECK(cudaMemsetAsync(lb.loss, 0, sizeof(float) * R, s));
    const uint32_t ray_offset = m_rank * R, n_rays_global = m_world * R;
    float* max_level = nullptr; float* max_level_compacted = nullptr;
    if (max_level_rand_training) {   // :3996: the inference pass sees each sample with its ray's level cut-off
        max_level = m_max_level.get<float>(max_samples); max_level_compacted = m_max_level_compacted.get<float>(target);
        nerf_fill_max_level(s, R, m_rng, m_n_images, m_metas.as<ImageMeta>(), train_snap_to_pixel_centers, ray_offset, n_rays_global, sb, max_level);
    }
    m_nerf_net->max_level_gpu = max_level;
    m_nerf_net->inference(s, m_trainer->params, sb.coords, max_inference, mlp_out);   // inference_mixed_precision(.., use_inference_params=false) (:4001)
    m_nerf_net->max_level_gpu = max_level_compacted;                                  // :4004, for the training pass below

    LossConfig lc{};
    lc.loss_type = loss_type; lc.rgb_activation = rgb_activation; lc.density_activation = density_activation;
    lc.random_bg_color = random_bg_color; lc.linear_colors = linear_colors; lc.snap_to_pixel_centers = train_snap_to_pixel_centers;
    memcpy(lc.background_color, background_color, sizeof(lc.background_color));
    lc.loss_scale = LOSS_SCALE; lc.near_distance = near_distance; lc.color_space = color_space;
    lc.max_level_rand = max_level_rand_training; lb.max_level_compacted = max_level_compacted;
    lc.depth_supervision_lambda = depth_supervision_lambda; lc.depth_loss_type = depth_loss_type;
    if (optimize_exposure) ensure_cam_exposure(s);
    if (m_cam_exposure_n == m_n_images && m_n_images) {   // trained (now or earlier): the targets carry the per-image exposure; gradients only while it is being optimised
        lc.exposure = m_cam_exposure.as<float>();
        if (optimize_exposure) {
            lc.exposure_gradient = m_cam_exposure_grad.as<float>();
            if (m_n_steps_since_cam_update == 0) NGP_CUDA_CHECK(cudaMemsetAsync(lc.exposure_gradient, 0, (size_t)m_n_images * 12, s));   // :3641-3644
        }
    }
    // error map (:3659-3666): sized for the rays one update window brings per view, cleared at the start of the window
    const bool error_sampling = sample_focal_plane_proportional_to_error || sample_image_proportional_to_error;
    if (error_sampling) {
        NGP_CHECK(!m_comm && m_world == 1, "error-map importance sampling is not available in data-parallel training");
        if (m_n_steps_since_error_map_update == 0 || !m_error_map_live) {
            const uint32_t n_samples_per_image = (uint32_t)(((uint64_t)n_steps_between_error_map_updates * R) / m_n_images);
            const int side = (int)(std::sqrt(std::sqrt((float)n_samples_per_image)) * 3.5f);
            m_error_map_res[0] = std::min(side, m_host_metas[0].res[0]); m_error_map_res[1] = std::min(side, m_host_metas[0].res[1]);
            const size_t n = (size_t)m_error_map_res[0] * m_error_map_res[1] * m_n_images;
            NGP_CUDA_CHECK(cudaMemsetAsync(m_error_map.get<float>(n), 0, n * sizeof(float), s));
            m_error_map_live = true; m_n_steps_since_error_map_update = 0;
        }
        lb.error_map = m_error_map.as<float>(); lb.error_map_res[0] = m_error_map_res[0]; lb.error_map_res[1] = m_error_map_res[1];
    } else m_error_map_live = false;
    float* dL_dcoords = nullptr;
    if (optimize_extrinsics || optimize_distortion) {   // train_camera (:4071-4073): the backward pass also returns dL/d(coords); accumulators cleared at the start of a window (:3641-3647)
        NGP_CHECK(!m_comm && m_world == 1, "optimize_extrinsics / optimize_distortion are not available in data-parallel training");
        dL_dcoords = m_coords_gradient.get<float>((size_t)target * 7);
    }
    if (optimize_extrinsics) {
        const bool fresh = m_cam_pos_grad.ptr == nullptr;
        float* gp = m_cam_pos_grad.get<float>((size_t)m_n_images * 3); float* gr = m_cam_rot_grad.get<float>((size_t)m_n_images * 3);
        if (fresh || m_n_steps_since_cam_update == 0) { NGP_CUDA_CHECK(cudaMemsetAsync(gp, 0, (size_t)m_n_images * 12, s)); NGP_CUDA_CHECK(cudaMemsetAsync(gr, 0, (size_t)m_n_images * 12, s)); }
    }
    if (optimize_distortion && m_n_steps_since_cam_update == 0) {   // (the map itself was created before the sampler ran: train_begin)
        const size_t n = (size_t)m_distortion_res[0] * m_distortion_res[1] * 2;
        NGP_CUDA_CHECK(cudaMemsetAsync(m_distortion_grad.as<float>(), 0, n * 4, s)); NGP_CUDA_CHECK(cudaMemsetAsync(m_distortion_weight.as<float>(), 0, n * 4, s));
    }
    lb.enc_in = m_nerf_net->last_inference_encoding();   // written by inference() just above
    if (m_copy_pending) { NGP_CUDA_CHECK(cudaStreamWaitEvent(s, m_ev_copy_done, 0)); m_copy_pending = false; }   // uploaded view(s), see set_training_image
    nerf_compute_loss(s, R, aabb, n_rays_total, m_rng, target, lc, m_n_images, m_metas.as<ImageMeta>(), m_mean.as<float>(), ray_offset, n_rays_global, sb, lb);
    NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counters, sb.counters, 16, cudaMemcpyDeviceToHost, s));   // read by train_end
    if (m_get_loss) {   // every 16th step (testbed.cu:4349-4353): the per-ray losses are final here, so the sum travels with the counters and nothing waits for the optimizer
        reduce_sum_f32(s, m_loss.as<float>(), R, m_loss_sum.as<float>(), m_partials.as<float>());
        NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counters + 4, m_loss_sum.as<float>(), 4, cudaMemcpyDeviceToHost, s));
        NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counters + 7, m_mean.as<float>(), 4, cudaMemcpyDeviceToHost, s));   // stats().density_grid_mean
    }
    NGP_CUDA_CHECK(cudaEventRecord(m_ev_counters, s));
    // device-side control law for the next step's sampler (single rank; data parallel keeps the host path: the law sees the rank mean)
    uint32_t early_rays_bound = 0, early_samples_bound = 0;
    {
        const uint32_t next_step = m_training_step + 1u, skip = std::min(std::max(next_step / 16u, 1u), 16u);
        const bool next_refreshes_grid = next_step % skip == 0;
        const uint32_t rb = std::min<uint32_t>(1u << 18, (uint32_t)next_multiple((uint64_t)R + R / 2 + BATCH_GRANULARITY, BATCH_GRANULARITY));
        const uint32_t lanes = sampler_lanes_per_ray ? sampler_lanes_per_ray : nerf_sampler_lanes_per_ray(rb);
        if (device_control_law && prefetch_rays && m_all_images_u8 && !m_comm && m_world == 1 && !next_refreshes_grid && lanes != 1 && !m_walk_ahead_valid && !optimize_extrinsics && !optimize_distortion && !error_sampling) {
            early_rays_bound = rb;
            early_samples_bound = std::min<uint32_t>(max_samples, (uint32_t)next_multiple((uint64_t)max_inference + max_inference / 2 + BATCH_GRANULARITY, BATCH_GRANULARITY));
            StepParams* sp = m_step_params[next_step % N_SETS].get<StepParams>(1);
            nerf_control_law(s, sb.counters, R, target, early_rays_bound, early_samples_bound, sp);
            NGP_CUDA_CHECK(cudaEventRecord(m_ev_after_loss, s));
        }
    }
    if (m_comm) {   // global sample counts for the control law: summed while the backward pass runs
        uint32_t* dc = m_dp_counters.as<uint32_t>();
        NGP_CUDA_CHECK(cudaMemcpyAsync(dc, sb.counters + 1, 8, cudaMemcpyDeviceToDevice, s));   // [1] before compaction, [2] compacted
        NGP_CUDA_CHECK(cudaEventRecord(m_ev_slice, s));
        NGP_CUDA_CHECK(cudaStreamWaitEvent(m_comm_stream, m_ev_slice, 0));
        m_comm->all_reduce_sum(dc, 2, COMM_U32, m_comm_stream);
        NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counters + 5, dc, 8, cudaMemcpyDeviceToHost, m_comm_stream));
        NGP_CUDA_CHECK(cudaEventRecord(m_ev_dp_counters, m_comm_stream));
    }
    nerf_fill_rollover(s, target, &sb.counters[2], lb.dloss_doutput, lb.coords_compacted, lb.enc_out, lb.enc_width, max_level_compacted);   // :4061-4069
    // forward(compacted) + backward(EGradientMode::Overwrite) (:4077-4078)
    if (m_comm && sharded_step()) {
        // sharded optimizer: ONE in-place reduce-scatter of the whole (padded) gradient vector on the communication stream; Adam on
        // this rank's slice and the parameter all-gather follow in train_end
        m_nerf_net->forward_backward_encoded(s, m_trainer->params, lb.coords_compacted, target, lb.dloss_doutput, m_trainer->grads, m_ev_after_mlp, m_trainer->grid_grads_clean);
        NGP_CUDA_CHECK(cudaEventRecord(m_ev_slice, s));
        NGP_CUDA_CHECK(cudaStreamWaitEvent(m_comm_stream, m_ev_slice, 0));
        m_comm->reduce_scatter_sum(m_trainer->grads, m_trainer->n_cap / (uint64_t)m_comm->world(), COMM_F16, m_comm_stream);
        NGP_CUDA_CHECK(cudaEventRecord(m_ev_comm_done, m_comm_stream));
        NGP_CUDA_CHECK(cudaStreamWaitEvent(s, m_ev_comm_done, 0));
    } else if (m_comm) {
        // every finished slice of the gradient vector (MLP weights, then one hash level chunk at a time) is summed over the ranks
        // on the communication stream while the next chunk is still being scattered; Adam waits for the last one
        __half* grads = m_trainer->grads;
        const std::function<void(size_t, size_t)> slice_ready = [&](size_t off, size_t count) {
            NGP_CUDA_CHECK(cudaEventRecord(m_ev_slice, s));
            NGP_CUDA_CHECK(cudaStreamWaitEvent(m_comm_stream, m_ev_slice, 0));
            m_comm->all_reduce_sum(grads + off, count, COMM_F16, m_comm_stream);
        };
        m_nerf_net->forward_backward_encoded(s, m_trainer->params, lb.coords_compacted, target, lb.dloss_doutput, grads, m_ev_after_mlp, m_trainer->grid_grads_clean, &slice_ready);
        NGP_CUDA_CHECK(cudaEventRecord(m_ev_comm_done, m_comm_stream));
        NGP_CUDA_CHECK(cudaStreamWaitEvent(s, m_ev_comm_done, 0));
    } else {
        m_nerf_net->forward_backward_encoded(s, m_trainer->params, lb.coords_compacted, target, lb.dloss_doutput, m_trainer->grads, m_ev_after_mlp, m_trainer->grid_grads_clean,
                                             nullptr, dL_dcoords);
        if (dL_dcoords)   // :4098-4124
            nerf_compute_cam_gradient(s, R, aabb, m_rng, train_snap_to_pixel_centers, m_metas.as<ImageMeta>(), m_n_images, ray_offset, n_rays_global, sb, lb.coords_compacted, dL_dcoords,
                                      optimize_extrinsics ? m_cam_pos_grad.as<float>() : nullptr, optimize_extrinsics ? m_cam_rot_grad.as<float>() : nullptr,
                                      optimize_distortion ? m_distortion_grad.as<float>() : nullptr, optimize_distortion ? m_distortion_weight.as<float>() : nullptr);
    }
    m_trainer->grid_grads_clean = false;
    m_nerf_net->max_level_gpu = nullptr;   // :4130
    m_rng.advance();   // :4127
    if (early_rays_bound) launch_early_prefetch(s, target, early_rays_bound, early_samples_bound);
}

// ---------------------------------------------------------------------------------------------------------------
// train_image (testbed_image.cu:214-285): Trainer::training_step = forward + L2 loss + backward
// ---------------------------------------------------------------------------------------------------------------
void Testbed::train_image_step(cudaStream_t s, uint32_t n, bool get_loss) {
    NGP_CHECK(m_image_w > 0, "Cannot train without training data");
    float* pos = m_img_pos.get<float>((size_t)n * 2); float* target = m_img_target.get<float>((size_t)n * 3);
    __half* out = m_img_out.get<__half>((size_t)n * 16); __half* dl = m_img_dl.get<__half>((size_t)n * 16);
    float* lossv = m_img_loss.get<float>(n); m_loss_sum.get<float>(4); m_partials.get<float>(1024);
    NGP_CHECK(image_random_mode >= IMG_RANDOM_RANDOM && image_random_mode <= IMG_RANDOM_STRATIFIED, "image.random_mode: unknown mode");
    image_generate_batch(s, n, m_rng, image_random_mode, (size_t)n * m_training_step, (uint32_t)seed, m_image.as<float>(), m_image_w, m_image_h,
                         image_snap_to_pixel_centers, image_linear_colors, pos, target);
    if (image_random_mode == IMG_RANDOM_RANDOM || image_random_mode == IMG_RANDOM_STRATIFIED) m_rng.advance();   // the low-discrepancy modes do not draw from m_rng
    m_image_net->inference(s, m_trainer->params, pos, 2, n, out);
    tcnn_loss(s, m_tcnn_loss, n, 3, out, target, LOSS_SCALE, dl, get_loss ? lossv : nullptr);
    if (get_loss) reduce_sum_f32(s, lossv, n, m_loss_sum.as<float>(), m_partials.as<float>());
    m_image_net->forward_backward(s, m_trainer->params, pos, 2, n, dl, m_trainer->grads);
}

// ---------------------------------------------------------------------------------------------------------------
// SDF mode with caller-supplied ground truth: override_sdf_training_data (python_api.cu:69-99) + train_sdf
// (testbed_sdf.cu:1289-1311).  Generating ground truth from a mesh (BVH / OptiX) is outside the path.
// ---------------------------------------------------------------------------------------------------------------
void Testbed::override_sdf_training_data(cudaStream_t s, const float* points, const float* distances, uint32_t n, const float* raw_aabb_min,
                                         const float* raw_aabb_max, float mesh_scale) {
    NGP_CHECK(m_mode == MODE_TB_SDF, "override_sdf_training_data: testbed is not in Sdf mode");
    NGP_CHECK(points && distances && n > 0, "Invalid Points<->Distances data");   // python_api.cu:73-76
    NGP_CHECK(mesh_scale > 0.f, "override_sdf_training_data: mesh_scale must be positive");
    // mesh space -> unit cube, as load_mesh leaves m_raw_aabb / mesh_scale (python_api.cu:85-86)
    std::vector<float> p((size_t)n * 3), d(n);
    for (uint32_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k)
            p[(size_t)i * 3 + k] = (points[(size_t)i * 3 + k] - raw_aabb_min[k]) / mesh_scale + 0.5f * (1.0f - (raw_aabb_max[k] - raw_aabb_min[k]) / mesh_scale);
        d[i] = distances[i] / mesh_scale;
    }
    NGP_CUDA_CHECK(cudaMemcpyAsync(m_sdf_pos.get<float>((size_t)n * 3), p.data(), (size_t)n * 12, cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaMemcpyAsync(m_sdf_dist.get<float>(n), d.data(), (size_t)n * 4, cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    m_sdf_size = n;
}

void Testbed::train_sdf_step(cudaStream_t s, uint32_t target_batch_size, bool get_loss) {
    NGP_CHECK(m_sdf_size > 0, "Cannot train without training data");
    m_sdf_step_skipped = m_sdf_size < target_batch_size;
    if (m_sdf_step_skipped) return;
    const uint32_t n = target_batch_size;   // min(size, target) with size >= target
    float* pos = m_img_pos.get<float>((size_t)n * 3); float* target = m_img_target.get<float>(n);
    __half* out = m_img_out.get<__half>((size_t)n * 16); __half* dl = m_img_dl.get<__half>((size_t)n * 16);
    float* lossv = m_img_loss.get<float>(n); m_loss_sum.get<float>(4); m_partials.get<float>(1024);
    // "Permute all training records to de-correlate training data": the first n records of permutation(seed = training step)
    sdf_shuffle(s, m_sdf_size, n, m_training_step, m_sdf_pos.as<float>(), m_sdf_dist.as<float>(), pos, target);
    m_image_net->inference(s, m_trainer->params, pos, 3, n, out);
    tcnn_loss(s, m_tcnn_loss, n, 1, out, target, LOSS_SCALE, dl, get_loss ? lossv : nullptr);
    if (get_loss) reduce_sum_f32(s, lossv, n, m_loss_sum.as<float>(), m_partials.as<float>());
    m_image_net->forward_backward(s, m_trainer->params, pos, 3, n, dl, m_trainer->grads);
}

// compute_image_mse (testbed_image.cu:465-527): all pixels at pixel centres, inference params, chunks of 2^20
float Testbed::compute_image_mse(cudaStream_t s, bool quantize) {
    NGP_CHECK(m_mode == MODE_TB_IMAGE && m_image_w > 0 && m_trainer, "compute_image_mse needs an image and a network");
    const uint32_t n_elements = (uint32_t)m_image_w * m_image_h, max_batch = 1u << 20;
    DeviceBuffer se_buf, pos_buf, tgt_buf, out_buf;
    float* se = se_buf.get<float>(n_elements + max_batch); float* pos = pos_buf.get<float>((size_t)max_batch * 2);
    float* tgt = tgt_buf.get<float>((size_t)max_batch * 3); __half* out = out_buf.get<__half>((size_t)max_batch * 16);
    m_loss_sum.get<float>(4); m_partials.get<float>(1024);
    for (uint32_t offset = 0; offset < n_elements; offset += max_batch) {
        const uint32_t batch = (std::min(max_batch, n_elements - offset) + 255u) & ~255u;
        image_eval_coords(s, batch, offset, m_image.as<float>(), m_image_w, m_image_h, image_linear_colors, pos, tgt);
        m_image_net->inference(s, m_trainer->inference_params, pos, 2, batch, out);
        image_squared_error(s, batch, tgt, out, se + offset, quantize);
    }
    reduce_sum_f32(s, se, n_elements, m_loss_sum.as<float>(), m_partials.as<float>());
    float sum = 0.f;
    NGP_CUDA_CHECK(cudaMemcpyAsync(&sum, m_loss_sum.as<float>(), 4, cudaMemcpyDeviceToHost, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    return sum / (float)n_elements;
}

// ---------------------------------------------------------------------------------------------------------------
// render_nerf + NerfTracer::trace (testbed_nerf.cu:2417-2693,2752-2871), Shade mode, headless float4 buffer
// ---------------------------------------------------------------------------------------------------------------
void Testbed::trace_sample(cudaStream_t s, const RenderCamera& cam, int sample, float* frame, float* fdepth) {
    NGP_CHECK(render_mode != RM_NORMALS, "render_mode Normals needs the network's input gradient: not implemented in this library");
    const uint32_t n_pixels = (uint32_t)cam.res[0] * (uint32_t)((cam.y1 - cam.y0 + cam.y_stride - 1) / cam.y_stride);
    const uint32_t n_pad = (uint32_t)next_multiple(n_pixels, BATCH_GRANULARITY);
    RaysSoa rays[3];
    for (int k = 0; k < 3; ++k) rays[k] = {m_r_rgba[k].get<float>((size_t)n_pad * 4), m_r_depth[k].get<float>(n_pad), m_r_payload[k].get<NerfPayload>(n_pad)};
    // rows the sample buffers hold.  An iteration asks for n_alive * clamp(target / n_alive, 1, cap) <= max(n_alive, target) rows, so max(n_pixels, target)
    // covers every case (the reference allocates n_pixels * MAX_STEPS rows, 2.4 GB at 3840 x 2160); small frames need no more than 32 steps per ray
    static const uint32_t target = [] { const char* e = getenv("NGP_TRACE_TARGET_SAMPLES"); return e ? (uint32_t)atol(e) : TRACE_TARGET_SAMPLES; }();
    const uint32_t capacity = (uint32_t)std::max<uint64_t>(n_pad, std::min<uint64_t>(target, (uint64_t)n_pad * TRACE_MAX_STEPS_TAIL));
    float* input = m_r_input.get<float>((size_t)capacity * 7); __half* output = m_r_output.get<__half>((size_t)capacity * 4);
    uint32_t* counters = m_r_counters.get<uint32_t>(4 + 2 * sizeof(TraceCtl) / 4); uint32_t* scan = m_r_scan.get<uint32_t>(div_round_up(n_pad, 256) + 1);
    TraceCtl* ctl = reinterpret_cast<TraceCtl*>(counters + 4);
    constexpr uint32_t RING = 8, LAG = 2;     // the host trails the device by LAG iterations when it reads the loop control back
    if (!m_trace_host) {
        NGP_CUDA_CHECK(cudaHostAlloc(reinterpret_cast<void**>(&m_trace_host), (RING + 1) * sizeof(TraceCtl), cudaHostAllocMapped));
        for (cudaEvent_t& e : m_ev_trace) NGP_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    const float depth_scale = 1.0f / m_ds_scale;   // render_nerf: 1 / dataset.scale
    // the coarse table speeds up walks that stay inside the unit cube at mip 0; rebuilt per call (4 KB) so that it always matches the bitfield
    const bool unit_cube = m_max_cascade == 0 && render_aabb.min[0] >= 0.f && render_aabb.min[1] >= 0.f && render_aabb.min[2] >= 0.f &&
                           render_aabb.max[0] <= 1.f && render_aabb.max[1] <= 1.f && render_aabb.max[2] <= 1.f;
    static const bool no_coarse = getenv("NGP_TRACE_NO_COARSE") != nullptr;
    const uint32_t* coarse = nullptr;
    if (unit_cube && !no_coarse) { uint32_t* c = m_r_coarse.get<uint32_t>(1024); nerf_build_coarse_occupancy(s, m_bitfield.as<uint8_t>(), c); coarse = c; }
    nerf_init_rays(s, cam, (uint32_t)sample, snap_to_pixel_centers, render_aabb, m_bitfield.as<uint8_t>(), m_max_cascade, cone_angle_constant, rays[0], frame, fdepth,
                   render_near_distance, coarse);
    // ctl[0] = state before the first compaction: every pixel is a ray, step index 1 (the reference's loop starts at i = 1)
    TraceCtl* boot = m_trace_host + RING;
    *boot = TraceCtl{n_pixels, 0, 0, 1};
    NGP_CUDA_CHECK(cudaMemsetAsync(counters, 0, 16, s));
    NGP_CUDA_CHECK(cudaMemcpyAsync(&ctl[0], boot, sizeof(TraceCtl), cudaMemcpyHostToDevice, s));
    uint32_t bound = n_pixels, db = 0;
    for (uint32_t it = 1;; ++it) {
        RaysSoa& cur = rays[(db + 1) % 2]; RaysSoa& tmp = rays[db % 2]; ++db;
        const TraceCtl* prev = &ctl[(it - 1) & 1]; TraceCtl* now = &ctl[it & 1];
        nerf_compact_rays(s, bound, tmp, cur, rays[2], counters, scan, prev, now, m_trace_host + it % RING, capacity, target);
        NGP_CUDA_CHECK(cudaEventRecord(m_ev_trace[it % RING], s));
        // rows of this iteration's network call are bounded by what `bound` rays can ask for; the kernels read the true counts from `now`
        const uint32_t rows_bound = (uint32_t)std::min<uint64_t>(capacity, next_multiple(std::max<uint64_t>(bound, std::min<uint64_t>(target, (uint64_t)bound * TRACE_MAX_STEPS_TAIL)), (uint64_t)BATCH_GRANULARITY));
        nerf_generate_next_inputs(s, bound, now, render_aabb, aabb, cur, input, m_bitfield.as<uint8_t>(), m_max_cascade, cone_angle_constant, coarse);
        m_nerf_net->inference(s, m_trainer->inference_params, input, rows_bound, output, &now->n_elements);
        nerf_composite(s, bound, now, aabb, cam.camera, cur, input, output, rgb_activation, density_activation, render_min_transmittance, render_mode, depth_scale);
        if (it >= LAG) {   // what iteration it - LAG + 1 left alive (long since computed unless the frame is tiny): tighter bounds, and the end
            const uint32_t seen = it - LAG + 1;
            NGP_CUDA_CHECK(cudaEventSynchronize(m_ev_trace[seen % RING]));
            const uint32_t alive = m_trace_host[seen % RING].n_alive;
            if (alive == 0) break;      // iterations seen + 1 .. it were enqueued with nothing to do and fall through
            bound = std::min(bound, alive);
        }
        NGP_CHECK(it < 10000, "render: ray marching did not terminate");   // MARCH_ITER (:100)
    }
    nerf_shade(s, n_pixels, counters + 1, rays[2], linear_colors, cam, frame, fdepth, render_mode);
}

void Testbed::render(cudaStream_t s, const float* camera_matrix, const float* focal_length, const float* screen_center_, int w, int h, int spp, int y0, int y1,
                     float* rgba_out, bool out_is_device, float* depth_out) {
    NGP_CHECK(m_mode == MODE_TB_NERF && m_nerf_net && m_trainer, "render: NeRF network not initialised");
    NGP_CHECK(w > 0 && h > 0 && y0 >= 0 && y1 <= h && y0 < y1 && spp >= 1 && render_row_stride >= 1, "render: bad resolution / row band");
    NGP_CHECK(!m_inference_stale, "render: sharded optimizer — call sync_inference_params (collective, every rank) first");
    const uint32_t n_pixels = (uint32_t)w * (uint32_t)((y1 - y0 + render_row_stride - 1) / render_row_stride);
    float* frame = m_frame.get<float>((size_t)n_pixels * 4); float* fdepth = m_frame_depth.get<float>(n_pixels); float* accum = m_accum.get<float>((size_t)n_pixels * 4);
    RenderCamera cam{};
    cam.res[0] = w; cam.res[1] = h; cam.y0 = y0; cam.y1 = y1; cam.y_stride = render_row_stride;
    memcpy(cam.focal_length, focal_length, 8); memcpy(cam.screen_center, screen_center_, 8); memcpy(cam.camera, camera_matrix, 48);
    cam.lens_mode = render_lens_mode; memcpy(cam.lens_params, render_lens_params, 16);
    NGP_CUDA_CHECK(cudaMemsetAsync(accum, 0, (size_t)n_pixels * 16, s));
    for (int sample = 0; sample < spp; ++sample) {
        trace_sample(s, cam, sample, frame, fdepth);
        frame_accumulate(s, n_pixels, frame, accum, (float)sample, CS_LINEAR);
    }
    NGP_CUDA_CHECK(cudaMemcpyAsync(rgba_out, accum, (size_t)n_pixels * 16, out_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    if (depth_out) NGP_CUDA_CHECK(cudaMemcpyAsync(depth_out, fdepth, (size_t)n_pixels * 4, out_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
}

// ---------------------------------------------------------------------------------------------------------------
// view state (testbed.cu:789-880,3810-3824) and render_to_cpu / render_frame with the output stage (python_api.cu:127-205,
// testbed.cu:4640-4760)
// ---------------------------------------------------------------------------------------------------------------
void Testbed::set_fov(float degrees) {   // fov_to_focal_length(1, degrees)
    relative_focal_length[0] = relative_focal_length[1] = 0.5f / tanf(0.5f * degrees * 3.14159265358979323846f / 180.0f);
}
float Testbed::fov() const {             // focal_length_to_fov(1, relative_focal_length[fov_axis])
    return 2.0f * 180.0f / 3.14159265358979323846f * atanf(1.0f / (relative_focal_length[fov_axis] * 2.0f));
}
void Testbed::reset_camera() {
    fov_axis = 1; zoom = 1.0f; screen_center[0] = screen_center[1] = 0.5f;
    if (m_mode == MODE_TB_IMAGE) { relative_focal_length[0] = relative_focal_length[1] = 1.0f; scale = 1.0f; }
    else { set_fov(50.625f); scale = 1.5f; }
    const float m[12] = {1, 0, 0, 0, -1, 0, 0, 0, -1, 0.5f, 0.5f, 0.5f};
    memcpy(camera, m, sizeof(m));
    for (int k = 0; k < 3; ++k) camera[9 + k] -= scale * camera[6 + k];
}
void Testbed::set_nerf_camera_matrix(const float* nerf) {   // NerfDataset::nerf_matrix_to_ngp (nerf_loader.h:120-140)
    float m[12]; memcpy(m, nerf, sizeof(m));
    for (int r = 0; r < 3; ++r) { m[3 + r] *= -1.f; m[6 + r] *= -1.f; m[9 + r] = m[9 + r] * m_ds_scale + m_ds_offset[r]; }
    for (int c = 0; c < 4; ++c) { camera[c * 3 + 0] = m[c * 3 + 1]; camera[c * 3 + 1] = m[c * 3 + 2]; camera[c * 3 + 2] = m[c * 3 + 0]; }
}
void Testbed::set_camera_to_training_view(int v) {
    NGP_CHECK(m_mode == MODE_TB_NERF && v >= 0 && v < (int)m_host_metas.size(), "set_camera_to_training_view: no such training view");   // views of a snapshot have poses but no pixels
    const ImageMeta& m = m_host_metas[v];
    float old_look_at[3]; for (int k = 0; k < 3; ++k) old_look_at[k] = camera[9 + k] + camera[6 + k] * scale;
    xform_given_rolling_shutter(m.xform, m.xform_end, m.rolling_shutter, 0.5f, 0.5f, 0.0f, camera);   // get_xform_given_rolling_shutter(.., {0.5, 0.5}, 0) (testbed.cu:843)
    for (int k = 0; k < 2; ++k) relative_focal_length[k] = m.focal_length[k] / (float)m.res[fov_axis];
    float d = 0.f; for (int k = 0; k < 3; ++k) d += (old_look_at[k] - camera[9 + k]) * camera[6 + k];
    scale = std::max(d, 0.1f);
    render_with_lens_distortion = true;
    render_lens_mode = m.lens_mode; memcpy(render_lens_params, m.lens_params, sizeof(render_lens_params));
    for (int k = 0; k < 2; ++k) screen_center[k] = 1.0f - m.principal_point[k];
    training_view = v;
}
void Testbed::apply_dataset_view_hints(bool has_render_aabb, const float* ra, const float* up3) {
    if (has_render_aabb) {   // m_render_aabb = dataset.render_aabb.intersection(m_aabb)
        for (int k = 0; k < 3; ++k) { render_aabb.min[k] = std::max(ra[k], aabb.min[k]); render_aabb.max[k] = std::min(ra[3 + k], aabb.max[k]); }
    }
    for (int k = 0; k < 3; ++k) { up_dir[k] = up3[k]; m_dataset_up[k] = up3[k]; }
}

void Testbed::render_frame(cudaStream_t s, int w, int h, int spp, bool to_srgb, int y0, int y1, float* rgba_out, bool out_is_device) {
    NGP_CHECK(m_mode == MODE_TB_NERF, "render: implemented for the Nerf mode");
    NGP_CHECK(w > 0 && h > 0 && y0 >= 0 && y1 <= h && y0 < y1 && spp >= 1 && render_row_stride >= 1, "render: bad resolution / row band");
    NGP_CHECK(!(render_ground_truth && render_row_stride != 1), "render_ground_truth: row_stride must be 1");
    NGP_CHECK(color_space >= CS_LINEAR && color_space <= CS_VISPOSNEG && tonemap_curve >= TM_IDENTITY && tonemap_curve <= TM_REINHARD, "render: bad color_space / tonemap_curve");
    const bool gt_only = render_ground_truth && ground_truth_alpha >= 1.0f;
    if (!gt_only) {
        NGP_CHECK(m_nerf_net && m_trainer, "render: NeRF network not initialised");
        NGP_CHECK(!m_inference_stale, "render: sharded optimizer — call sync_inference_params (collective, every rank) first");
    }
    const uint32_t n_pixels = (uint32_t)w * (uint32_t)((y1 - y0 + render_row_stride - 1) / render_row_stride);
    float* frame = m_frame.get<float>((size_t)n_pixels * 4); float* fdepth = m_frame_depth.get<float>(n_pixels); float* accum = m_accum.get<float>((size_t)n_pixels * 4);
    float* surface = m_surface.get<float>((size_t)n_pixels * 4);
    RenderCamera cam{};
    cam.res[0] = w; cam.res[1] = h; cam.y0 = y0; cam.y1 = y1; cam.y_stride = render_row_stride;
    const int res[2] = {w, h};
    for (int k = 0; k < 2; ++k) {
        cam.focal_length[k] = relative_focal_length[k] * (float)res[fov_axis] * zoom;     // calc_focal_length
        cam.screen_center[k] = (0.5f - screen_center[k]) * zoom + 0.5f;                    // render_screen_center
    }
    memcpy(cam.camera, camera, sizeof(camera));
    if (render_with_lens_distortion) { cam.lens_mode = render_lens_mode; memcpy(cam.lens_params, render_lens_params, 16); cam.distortion = distortion_view(); }   // testbed_nerf.cu:2776-2779
    const int out_cs = to_srgb ? CS_SRGB : CS_LINEAR;
    NGP_CUDA_CHECK(cudaMemsetAsync(accum, 0, (size_t)n_pixels * 16, s));
    for (int sample = 0; sample < spp; ++sample) {
        if (gt_only) NGP_CUDA_CHECK(cudaMemsetAsync(frame, 0, (size_t)n_pixels * 16, s));   // the frame buffer is cleared, nothing is traced
        else trace_sample(s, cam, sample, frame, fdepth);
        frame_accumulate(s, n_pixels, frame, accum, (float)sample, color_space);
    }
    frame_tonemap(s, n_pixels, exposure, background_color, accum, color_space, out_cs, tonemap_curve, false, false, surface);
    if (render_ground_truth) {
        NGP_CHECK(training_view >= 0 && training_view < (int)m_n_images, "render_ground_truth: no training view selected");
        const ImageMeta& m = m_host_metas[training_view];
        NGP_CHECK(m.pixels, "render_ground_truth: the training view has no pixels (metadata restored from a snapshot)");
        float e3[3] = {exposure, exposure, exposure};   // + cam_exposure[view] (testbed.cu:4726)
        if (m_cam_exposure_n == m_n_images && m_n_images) {
            float ce[3];
            NGP_CUDA_CHECK(cudaMemcpyAsync(ce, m_cam_exposure.as<float>() + (size_t)training_view * 3, 12, cudaMemcpyDeviceToHost, s));
            NGP_CUDA_CHECK(cudaStreamSynchronize(s));
            for (int k = 0; k < 3; ++k) e3[k] += ce[k];
        }
        const float centre[2] = {0.5f, 0.5f};
        frame_overlay_image(s, w, h, y0, y1, ground_truth_alpha, e3, background_color, out_cs, m.pixels, m.data_type, m.res, tonemap_curve, color_space, fov_axis, zoom,
                            centre, surface);
    }
    NGP_CUDA_CHECK(cudaMemcpyAsync(rgba_out, surface, (size_t)n_pixels * 16, out_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
}


void Testbed::debug_buffer(const std::string& name, void** ptr, uint64_t* bytes) {
    struct E { const char* n; DeviceBuffer* b; };
    const uint32_t k = m_last_set;   // the buffer set the last training step used (the other one may hold the next step's prefetch)
    const E table[] = {{"coords", &m_coords[k]}, {"numsteps", &m_numsteps[k]}, {"ray_indices", &m_ray_indices[k]}, {"rays", &m_rays[k]}, {"counters", &m_counters[k]},
                       {"mlp_out", &m_mlp_out}, {"dloss", &m_dloss}, {"coords_compacted", &m_coords_compacted}, {"loss", &m_loss},
                       {"grid_pos", &m_grid_pos}, {"grid_idx", &m_grid_idx}, {"grid_out", &m_grid_out}, {"density_grid_tmp", &m_density_grid_tmp},
                       {"mean_density", &m_mean}, {"cam_exposure", &m_cam_exposure}, {"cam_exposure_gradient", &m_cam_exposure_grad}, {"numsteps_all", &m_numsteps_all[k]}, {"image_pos", &m_img_pos}, {"image_target", &m_img_target},
                       {"image_out", &m_img_out}, {"image_dl", &m_img_dl}, {"image", &m_image}, {"max_level", &m_max_level}, {"max_level_compacted", &m_max_level_compacted}, {"distortion_map", &m_distortion_map}, {"distortion_gradient", &m_distortion_grad}, {"distortion_gradient_weight", &m_distortion_weight}, {"error_map", &m_error_map}, {"cdf_x_cond_y", &m_cdf_x_cond_y}, {"cdf_y", &m_cdf_y}, {"cdf_img", &m_cdf_img}, {"cam_pos_gradient", &m_cam_pos_grad}, {"cam_rot_gradient", &m_cam_rot_grad}, {"coords_gradient", &m_coords_gradient}, {"sdf_positions", &m_sdf_pos}, {"sdf_distances", &m_sdf_dist}};
    for (const E& e : table) if (name == e.n) { *ptr = e.b->ptr; *bytes = e.b->bytes; return; }
    throw Error("debug_buffer: unknown buffer '" + name + "'");
}

}  // namespace ngp
