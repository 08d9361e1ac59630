// Snapshot (.ingp / .msgpack) write and read: Testbed::save_snapshot / load_snapshot of the reference
// (src/testbed.cu:4873-4936, :4938-5057; load_network_config :246-273).  A snapshot is the network config with a "snapshot"
// object added, serialised as MessagePack; ".ingp" files are the same bytes inside a zlib stream (zstr::ostream / istream).
// Trainer::serialize / deserialize and the optimizers' serialize() are tiny-cuda-nn code (absent from the mount): key names
// are restated from its published source ("n_params", "params_type", "params_binary" = the fp16 inference parameters;
// optimizer: Ema{"nested"}, ExponentialDecay{"nested"}, Adam{"current_step", "base_learning_rate", "first_moments_binary",
// "second_moments_binary", "param_steps_binary"}) — unpinned until a stock snapshot is available.
#include "testbed.h"
#include "msgpack_mini.h"
#include <zlib.h>
#include <algorithm>
#include <cstdio>
#include <fstream>

namespace ngp {

static const int SNAPSHOT_FORMAT_VERSION = 1;   // testbed.cu:80

__global__ void half_to_float_kernel(uint64_t n, const __half* __restrict__ in, float* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __half2float(in[i]);
}
__global__ void float_to_half_kernel(uint64_t n, const float* __restrict__ in, __half* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2half_rn(in[i]);
}
static void half_to_float(cudaStream_t s, uint64_t n, const __half* in, float* out) {
    if (!n) return;
    half_to_float_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, s>>>(n, in, out); NGP_POST_LAUNCH();
}
static void float_to_half(cudaStream_t s, uint64_t n, const float* in, __half* out) {
    if (!n) return;
    float_to_half_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, s>>>(n, in, out); NGP_POST_LAUNCH();
}

static Json device_binary(cudaStream_t s, const void* dev, size_t bytes) {   // to_json(GPUMemory<T>): the raw bytes as msgpack bin
    Json j; j.type = Json::Binary; j.str.resize(bytes);
    if (bytes) NGP_CUDA_CHECK(cudaMemcpyAsync(&j.str[0], dev, bytes, cudaMemcpyDeviceToHost, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    return j;
}
static void binary_to_device(cudaStream_t s, const Json& j, void* dev, size_t bytes, const char* what) {
    NGP_CHECK(j.type == Json::Binary && j.str.size() == bytes, std::string("snapshot: '") + what + "' has the wrong size");
    if (bytes) NGP_CUDA_CHECK(cudaMemcpyAsync(dev, j.str.data(), bytes, cudaMemcpyHostToDevice, s));
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
}

static Json vec_json(const float* v, int n) { return Json::array_of(v, (size_t)n); }
static Json aabb_json(const Aabb& b) { Json j; j["min"] = vec_json(b.min, 3); j["max"] = vec_json(b.max, 3); return j; }
static void aabb_from_json(const Json& j, Aabb& b) {
    for (int k = 0; k < 3; ++k) { b.min[k] = (float)j.at("min").arr.at(k).num; b.max[k] = (float)j.at("max").arr.at(k).num; }
}
// glm mat4x3 (column major, 12 floats) <-> 3 rows of 4 (json_binding.h:45-66)
static Json mat4x3_json(const float* m) {
    Json j; j.type = Json::Array;
    for (int row = 0; row < 3; ++row) { Json r; r.type = Json::Array; for (int col = 0; col < 4; ++col) r.arr.push_back(Json::number(m[col * 3 + row])); j.arr.push_back(r); }
    return j;
}
static void mat4x3_from_json(const Json& j, float* m) {
    for (int row = 0; row < 3; ++row) for (int col = 0; col < 4; ++col) m[col * 3 + row] = (float)j.arr.at(row).arr.at(col).num;
}

static Json lens_json(const ImageMeta& m) {   // to_json(Lens) (json_binding.h:162-187)
    Json j; j.type = Json::Object;
    if (m.lens_mode == LENS_OPENCV) {
        j["is_fisheye"] = Json::boolean(false);
        j["k1"] = Json::number(m.lens_params[0]); j["k2"] = Json::number(m.lens_params[1]);
        j["p1"] = Json::number(m.lens_params[2]); j["p2"] = Json::number(m.lens_params[3]);
    }
    return j;
}

// to_json(NerfDataset) (json_binding.h:224-251): poses and intrinsics of the training views, no pixels
static Json dataset_json(const std::vector<ImageMeta>& metas, const Aabb& render_aabb, int aabb_scale, bool is_hdr, float scale, const float* offset, const float* up) {
    Json j; j.type = Json::Object;
    j["n_images"] = Json::integer((int64_t)metas.size());
    Json paths, md, xf; paths.type = md.type = xf.type = Json::Array;
    for (const ImageMeta& m : metas) {
        paths.arr.push_back(Json::string(""));
        Json e;
        e["focal_length"] = vec_json(m.focal_length, 2);
        e["lens"] = lens_json(m);
        e["principal_point"] = vec_json(m.principal_point, 2);
        e["rolling_shutter"] = vec_json(m.rolling_shutter, 4);
        Json res; res.type = Json::Array; res.arr.push_back(Json::integer(m.res[0])); res.arr.push_back(Json::integer(m.res[1]));
        e["resolution"] = res;
        md.arr.push_back(e);
        Json x; x["start"] = mat4x3_json(m.xform); x["end"] = mat4x3_json(m.xform_end);
        xf.arr.push_back(x);
    }
    j["paths"] = paths; j["metadata"] = md; j["xforms"] = xf;
    j["render_aabb"] = aabb_json(render_aabb);
    const float eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    Json r2l; r2l.type = Json::Array; for (int r = 0; r < 3; ++r) r2l.arr.push_back(vec_json(eye + 3 * r, 3));
    j["render_aabb_to_local"] = r2l;
    j["up"] = vec_json(up, 3); j["offset"] = vec_json(offset, 3);          // the loaded dataset's scale / offset (nerf_loader.cu:388-393) and up vector (what load_nerf_post copied into the testbed's)
    Json er; er.type = Json::Array; er.arr.push_back(Json::integer(0)); er.arr.push_back(Json::integer(0));
    j["envmap_resolution"] = er;
    j["scale"] = Json::number(scale);
    j["aabb_scale"] = Json::integer(aabb_scale);
    j["from_mitsuba"] = Json::boolean(false);
    j["is_hdr"] = Json::boolean(is_hdr);
    j["wants_importance_sampling"] = Json::boolean(false);
    j["n_extra_learnable_dims"] = Json::integer(0);
    return j;
}

static const char* mode_string(int mode) { return mode == MODE_TB_NERF ? "nerf" : mode == MODE_TB_SDF ? "sdf" : "image"; }   // common.cu:175-184

static bool ends_with_ci(const std::string& s, const char* ext) {
    const std::string e = ext;
    return s.size() >= e.size() && to_lower(s.substr(s.size() - e.size())) == e;
}

void Testbed::save_snapshot(cudaStream_t s, const std::string& path, bool include_optimizer_state, bool compress) {
    NGP_CHECK(m_trainer, "save_snapshot: no network to save");
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));
    // Trainer::serialize(include_optimizer_state)
    Json snap; snap.type = Json::Object;
    const uint64_t n = m_trainer->n_params;
    snap["n_params"] = Json::integer((int64_t)n);
    snap["params_type"] = Json::string("__half");
    snap["params_binary"] = device_binary(s, m_trainer->inference_params, n * 2);
    if (include_optimizer_state) {
        Json adam; adam.type = Json::Object;
        adam["current_step"] = Json::integer(m_trainer->opt.step_count);
        adam["base_learning_rate"] = Json::number(m_trainer->opt.learning_rate);
        adam["first_moments_binary"] = device_binary(s, m_trainer->m, n * 4);
        adam["second_moments_binary"] = device_binary(s, m_trainer->v, n * 4);
        adam["param_steps_binary"] = device_binary(s, m_trainer->steps, n * 4);
        // not in tcnn's Adam: the fp32 master weights and (under Ema) the running average, so that training resumes exactly
        adam["params_full_precision_binary"] = device_binary(s, m_trainer->master, n * 4);
        adam["current_learning_rate"] = Json::number(m_trainer->opt.current_lr);
        Json o = adam;
        if (m_trainer->opt.decay_interval > 0 || m_trainer->opt.decay_base != 1.0f) { Json d; d["nested"] = o; o = d; }
        if (m_trainer->ema) { Json e; e["nested"] = o; e["ema_binary"] = device_binary(s, m_trainer->ema, n * 4); o = e; }
        snap["optimizer"] = o;
    }
    snap["version"] = Json::integer(SNAPSHOT_FORMAT_VERSION);
    snap["mode"] = Json::string(mode_string(m_mode));
    if (m_mode == MODE_TB_NERF) {
        snap["density_grid_size"] = Json::integer(NERF_GRIDSIZE);
        const size_t cells = m_grid_allocated ? (size_t)NERF_GRID_N_CELLS * (m_max_cascade + 1) : 0;
        DeviceBuffer h16;
        if (cells) float_to_half(s, cells, m_density_grid.as<float>(), h16.get<__half>(cells));
        snap["density_grid_binary"] = device_binary(s, h16.ptr, cells * 2);
        Json nerf; nerf.type = Json::Object;
        nerf["aabb_scale"] = Json::integer(m_aabb_scale);
        Json rgb;
        rgb["rays_per_batch"] = Json::integer(m_rays_per_batch);
        rgb["measured_batch_size"] = Json::integer(m_measured);
        rgb["measured_batch_size_before_compaction"] = Json::integer(m_measured_before);
        nerf["rgb"] = rgb;
        // the data set keeps the poses as given (NerfDataset::xforms); the trained offsets travel beside them (testbed.cu:4893-4894) and are
        // applied again by update_transforms when the snapshot is loaded (testbed.cu:5010-5013)
        std::vector<ImageMeta> given = m_host_metas;
        for (size_t i = 0; i < given.size() && i < m_base_xforms.size(); ++i) { memcpy(given[i].xform, m_base_xforms[i].data(), 48); memcpy(given[i].xform_end, m_base_xforms[i].data() + 12, 48); }
        nerf["dataset"] = dataset_json(given, render_aabb, m_aabb_scale, !m_all_images_u8, m_ds_scale, m_ds_offset, m_dataset_up);
        const auto opt_json = [](const std::vector<VecAdam>& v, float lr) {   // AdamOptimizer<vec3> / RotationAdamOptimizer::to_json (adam_optimizer.h:177-186,265-274)
            Json a; a.type = Json::Array;
            for (const VecAdam& o : v) {
                Json e; e["iter"] = Json::integer(o.iter); e["first_moment"] = vec_json(o.m1, 3); e["second_moment"] = vec_json(o.m2, 3); e["variable"] = vec_json(o.var, 3);
                e["learning_rate"] = Json::number(lr); e["epsilon"] = Json::number(1e-8f); e["beta1"] = Json::number(0.9f); e["beta2"] = Json::number(0.99f);
                a.arr.push_back(e);
            }
            return a;
        };
        nerf["cam_pos_offset"] = opt_json(m_cam_pos_offset, 1e-4f); nerf["cam_rot_offset"] = opt_json(m_cam_rot_offset, 1e-4f);
        snap["nerf"] = nerf;
    }
    snap["training_step"] = Json::integer(m_training_step);
    snap["loss"] = Json::number(m_loss_scalar);
    snap["aabb"] = aabb_json(aabb);
    snap["render_aabb"] = aabb_json(render_aabb);
    snap["background_color"] = vec_json(background_color, 4);
    // the testbed's own view state (testbed.cu:4903-4913)
    snap["up_dir"] = vec_json(up_dir, 3);
    snap["exposure"] = Json::number(exposure);
    {
        Json cam; cam["matrix"] = mat4x3_json(camera); cam["fov_axis"] = Json::integer(fov_axis); cam["relative_focal_length"] = vec_json(relative_focal_length, 2);
        cam["screen_center"] = vec_json(screen_center, 2); cam["zoom"] = Json::number(zoom); cam["scale"] = Json::number(scale);
        snap["camera"] = cam;
    }
    Json cfg = m_config; cfg["snapshot"] = snap;

    std::string bytes;
    msgpack_encode(cfg, bytes);
    if (ends_with_ci(path, ".ingp")) {   // zlib stream, as zstr::ostream writes it (testbed.cu:4926-4929)
        uLongf cap = compressBound((uLong)bytes.size());
        NGP_CHECK(bytes.size() < 0xFFFFFFFFull, "save_snapshot: snapshot larger than 4 GiB");
        std::string z(cap, '\0');
        const int rc = compress2((Bytef*)&z[0], &cap, (const Bytef*)bytes.data(), (uLong)bytes.size(), compress ? Z_DEFAULT_COMPRESSION : Z_NO_COMPRESSION);
        NGP_CHECK(rc == Z_OK, "save_snapshot: zlib compression failed");
        z.resize(cap); bytes.swap(z);
    }
    std::ofstream f(path, std::ios::out | std::ios::binary);
    NGP_CHECK((bool)f, "save_snapshot: cannot open '" + path + "' for writing");
    f.write(bytes.data(), (std::streamsize)bytes.size());
    NGP_CHECK((bool)f, "save_snapshot: write to '" + path + "' failed");
}

static std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::in | std::ios::binary);
    if (!f) throw Error("Network snapshot '" + path + "' does not exist.");   // testbed.cu:249-251
    return std::string((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

static std::string zlib_inflate(const std::string& in) {   // zstr::istream: zlib or gzip framing, detected from the header
    z_stream z; memset(&z, 0, sizeof(z));
    NGP_CHECK(inflateInit2(&z, 15 + 32) == Z_OK, "load_snapshot: zlib init failed");
    std::string out; std::vector<char> buf(1 << 20);
    z.next_in = (Bytef*)in.data(); z.avail_in = (uInt)std::min<size_t>(in.size(), 0xFFFFFFFFu);
    int rc = Z_OK;
    while (rc != Z_STREAM_END) {
        z.next_out = (Bytef*)buf.data(); z.avail_out = (uInt)buf.size();
        rc = inflate(&z, Z_NO_FLUSH);
        if (rc != Z_OK && rc != Z_STREAM_END) { inflateEnd(&z); throw Error("load_snapshot: corrupt zlib stream in .ingp file"); }
        out.append(buf.data(), buf.size() - z.avail_out);
        if (rc == Z_OK && z.avail_in == 0 && z.avail_out != 0) { inflateEnd(&z); throw Error("load_snapshot: truncated .ingp file"); }
    }
    inflateEnd(&z);
    return out;
}

// Adam leaf of a serialised optimizer chain: Ema{nested} / ExponentialDecay{nested} / Adam
static const Json* find_adam(const Json& o, const Json** ema_level) {
    const Json* cur = &o;
    for (int depth = 0; depth < 4 && cur->is_object(); ++depth) {
        if (cur->contains("ema_binary") && ema_level) *ema_level = cur;
        if (cur->contains("first_moments_binary")) return cur;
        if (!cur->contains("nested")) break;
        cur = &cur->at("nested");
    }
    return nullptr;
}

int Testbed::snapshot_mode(const std::string& path) {
    std::string bytes = read_file(path);
    if (ends_with_ci(path, ".ingp")) bytes = zlib_inflate(bytes);
    Json config = msgpack_decode(bytes.data(), bytes.size());
    if (!config.contains("snapshot")) throw Error("File '" + path + "' does not contain a snapshot.");
    const Json& snap = config.at("snapshot");
    if (!snap.contains("mode")) return MODE_TB_NERF;
    const std::string mode = to_lower(snap.at("mode").str);
    if (mode == "nerf") return MODE_TB_NERF;
    if (mode == "sdf") return MODE_TB_SDF;
    if (mode == "image") return MODE_TB_IMAGE;
    if (mode == "volume") return 3;                                  // ETestbedMode::Volume: a mode this library does not implement
    return 4;                                                        // ETestbedMode::None
}

void Testbed::load_snapshot(cudaStream_t s, const std::string& path) {
    std::string bytes = read_file(path);
    if (ends_with_ci(path, ".ingp")) bytes = zlib_inflate(bytes);
    Json config = msgpack_decode(bytes.data(), bytes.size());
    bytes.clear(); bytes.shrink_to_fit();
    if (!config.contains("snapshot")) throw Error("File '" + path + "' does not contain a snapshot.");
    const Json& snap = config.at("snapshot");
    if (snap.value("version", 0.0) < SNAPSHOT_FORMAT_VERSION) throw Error("Snapshot uses an old format and can not be loaded.");
    const std::string mode = snap.contains("mode") ? to_lower(snap.at("mode").str) : (snap.contains("nerf") ? "nerf" : "");
    NGP_CHECK(mode == mode_string(m_mode), "load_snapshot: snapshot mode '" + mode + "' does not match this testbed's mode '" + mode_string(m_mode) + "'");
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
    if (m_aux_stream) NGP_CUDA_CHECK(cudaStreamSynchronize(m_aux_stream));
    m_prefetch_valid = false;

    bool dataset_from_snapshot = false;
    if (m_mode == MODE_TB_NERF) {
        if ((uint32_t)snap.value("density_grid_size", 0.0) != NERF_GRIDSIZE) throw Error("Incompatible grid size.");
        const Json& nerf = snap.at("nerf");
        int aabb_scale = m_aabb_scale;
        if (nerf.contains("aabb_scale")) aabb_scale = (int)nerf.at("aabb_scale").num;
        if (m_n_images == 0 && nerf.contains("dataset")) {
            // no data set loaded: take the views' poses and intrinsics from the snapshot (render only, no pixels; testbed.cu:4976-4979)
            const Json& d = nerf.at("dataset");
            const size_t n_img = (size_t)d.value("n_images", 0.0);
            m_host_metas.assign(n_img, ImageMeta{});
            for (size_t i = 0; i < n_img; ++i) {
                ImageMeta& m = m_host_metas[i];
                m.pixels = nullptr; m.data_type = IMG_BYTE;
                const Json& md = d.at("metadata").arr.at(i);
                for (int k = 0; k < 2; ++k) {
                    m.focal_length[k] = (float)md.at("focal_length").arr.at(k).num; m.principal_point[k] = (float)md.at("principal_point").arr.at(k).num;
                    m.res[k] = (int)md.at("resolution").arr.at(k).num;
                }
                m.lens_mode = LENS_PERSPECTIVE;
                if (md.contains("lens") && md.at("lens").contains("k1") && !md.at("lens").value("is_fisheye", false)) {
                    m.lens_mode = LENS_OPENCV; const char* keys[4] = {"k1", "k2", "p1", "p2"};
                    for (int k = 0; k < 4; ++k) m.lens_params[k] = (float)md.at("lens").value(keys[k], 0.0);
                }
                mat4x3_from_json(d.at("xforms").arr.at(i).at("start"), m.xform);
                if (d.at("xforms").arr.at(i).contains("end")) mat4x3_from_json(d.at("xforms").arr.at(i).at("end"), m.xform_end); else memcpy(m.xform_end, m.xform, sizeof(m.xform));
                if (md.contains("rolling_shutter") && md.at("rolling_shutter").arr.size() == 4) for (int k = 0; k < 4; ++k) m.rolling_shutter[k] = (float)md.at("rolling_shutter").arr[k].num;
            }
            m_base_xforms.resize(n_img);
            for (size_t i = 0; i < n_img; ++i) { memcpy(m_base_xforms[i].data(), m_host_metas[i].xform, 48); memcpy(m_base_xforms[i].data() + 12, m_host_metas[i].xform_end, 48); }
            m_cam_pos_offset.assign(n_img, VecAdam{}); m_cam_rot_offset.assign(n_img, VecAdam{});
            dataset_from_snapshot = true;
            NGP_CHECK(d.value("n_extra_learnable_dims", 0.0) <= 0.0, "load_snapshot: a network with per-view latent codes (n_extra_learnable_dims > 0) is not supported");
            aabb_scale = (int)d.value("aabb_scale", (double)aabb_scale);
            m_ds_scale = (float)d.value("scale", (double)m_ds_scale);
            if (d.contains("offset") && d.at("offset").arr.size() == 3) for (int k = 0; k < 3; ++k) m_ds_offset[k] = (float)d.at("offset").arr[k].num;
            if (d.contains("up") && d.at("up").arr.size() == 3) for (int k = 0; k < 3; ++k) m_dataset_up[k] = (float)d.at("up").arr[k].num;
        }
        load_nerf_post(s, aabb_scale);
        m_rays_per_batch = (uint32_t)nerf.at("rgb").at("rays_per_batch").num;
        m_measured = (uint32_t)nerf.at("rgb").at("measured_batch_size").num;
        m_measured_before = (uint32_t)nerf.at("rgb").at("measured_batch_size_before_compaction").num;
    }
    if (snap.contains("aabb")) aabb_from_json(snap.at("aabb"), aabb);
    if (snap.contains("render_aabb")) aabb_from_json(snap.at("render_aabb"), render_aabb);
    if (snap.contains("background_color")) for (size_t k = 0; k < std::min<size_t>(4, snap.at("background_color").arr.size()); ++k) background_color[k] = (float)snap.at("background_color").arr.at(k).num;
    // the view state (testbed.cu:5006-5031; "Needs to happen after load_nerf_post()")
    const auto read_vec = [](const Json& j, float* out, size_t n) { if (j.type == Json::Array && j.arr.size() == n) for (size_t k = 0; k < n; ++k) out[k] = (float)j.arr[k].num; };
    exposure = (float)snap.value("exposure", (double)exposure);
    if (snap.contains("up_dir")) read_vec(snap.at("up_dir"), up_dir, 3);
    if (snap.contains("camera")) {
        const Json& cam = snap.at("camera");
        if (cam.contains("matrix")) mat4x3_from_json(cam.at("matrix"), camera);
        fov_axis = (int)cam.value("fov_axis", (double)fov_axis);
        if (cam.contains("relative_focal_length")) read_vec(cam.at("relative_focal_length"), relative_focal_length, 2);
        if (cam.contains("screen_center")) read_vec(cam.at("screen_center"), screen_center, 2);
        zoom = (float)cam.value("zoom", (double)zoom); scale = (float)cam.value("scale", (double)scale);
    }

    // m_network_config = config; reset_network(false); then counters, Trainer::deserialize (testbed.cu:5028-5040)
    const uint32_t rays_per_batch = m_rays_per_batch, measured = m_measured, measured_before = m_measured_before;
    m_config = config; m_config.obj.erase("snapshot"); m_have_config = true;
    reset_network();
    m_rays_per_batch = rays_per_batch; m_measured = measured; m_measured_before = measured_before;
    m_training_step = (uint32_t)snap.at("training_step").num;
    m_loss_scalar = (float)snap.at("loss").num;

    const uint64_t n = m_trainer->n_params;
    NGP_CHECK((uint64_t)snap.value("n_params", 0.0) == n, "load_snapshot: the snapshot's parameter count does not match its network config");
    const std::string ptype = snap.value("params_type", "__half");
    if (ptype == "__half") {   // Trainer::set_params: the blob becomes params, inference params and (converted) the fp32 copy
        binary_to_device(s, snap.at("params_binary"), m_trainer->params, n * 2, "params_binary");
        half_to_float(s, n, m_trainer->params, m_trainer->master);
    } else if (ptype == "float") {
        binary_to_device(s, snap.at("params_binary"), m_trainer->master, n * 4, "params_binary");
        float_to_half(s, n, m_trainer->master, m_trainer->params);
    } else throw Error("Snapshot parameters must be of type float of __half");
    if (m_trainer->inference_params != m_trainer->params)
        NGP_CUDA_CHECK(cudaMemcpyAsync(m_trainer->inference_params, m_trainer->params, n * 2, cudaMemcpyDeviceToDevice, s));
    if (m_trainer->ema) half_to_float(s, n, m_trainer->params, m_trainer->ema);   // the running average restarts from the loaded weights
    m_trainer->opt.step_count = m_training_step;
    if (snap.contains("optimizer")) {
        const Json* ema_level = nullptr;
        const Json* adam = find_adam(snap.at("optimizer"), &ema_level);
        NGP_CHECK(adam, "load_snapshot: optimizer state has no Adam leaf");
        binary_to_device(s, adam->at("first_moments_binary"), m_trainer->m, n * 4, "first_moments_binary");
        binary_to_device(s, adam->at("second_moments_binary"), m_trainer->v, n * 4, "second_moments_binary");
        binary_to_device(s, adam->at("param_steps_binary"), m_trainer->steps, n * 4, "param_steps_binary");
        m_trainer->opt.step_count = (uint32_t)adam->value("current_step", (double)m_training_step);
        if (adam->contains("params_full_precision_binary")) {
            binary_to_device(s, adam->at("params_full_precision_binary"), m_trainer->master, n * 4, "params_full_precision_binary");
            float_to_half(s, n, m_trainer->master, m_trainer->params);
        }
        if (adam->contains("current_learning_rate")) m_trainer->opt.current_lr = (float)adam->at("current_learning_rate").num;
        if (ema_level && m_trainer->ema) binary_to_device(s, ema_level->at("ema_binary"), m_trainer->ema, n * 4, "ema_binary");
    }
    // the learning-rate schedule is a function of the step count: replay it when the snapshot carries no current rate
    if (!(snap.contains("optimizer") && find_adam(snap.at("optimizer"), nullptr)->contains("current_learning_rate"))) {
        Optimizer& o = m_trainer->opt; o.current_lr = o.learning_rate;
        if (o.decay_interval > 0)
            for (uint32_t sc = o.decay_start; sc <= o.step_count && sc < o.decay_end; sc += o.decay_interval) if (sc >= 1) o.current_lr *= o.decay_base;
    }

    if (m_mode == MODE_TB_NERF && snap.at("nerf").contains("dataset")) {
        // "If the snapshot appears to come from the same dataset as was already present (or none was previously present ...), load dataset-specific
        // optimized quantities" (testbed.cu:5043-5054): the per-view pose offsets; the poses the kernels see are rebuilt from them
        const Json& nerf = snap.at("nerf"); const Json& d = nerf.at("dataset");
        bool same = dataset_from_snapshot;
        if (!same && (size_t)d.value("n_images", 0.0) == m_base_xforms.size() && d.contains("xforms") && d.at("xforms").arr.size() == m_base_xforms.size()) {
            same = true;
            for (size_t i = 0; i < m_base_xforms.size() && same; ++i) { float x[12]; mat4x3_from_json(d.at("xforms").arr[i].at("start"), x); same = memcmp(x, m_base_xforms[i].data(), 48) == 0; }
        }
        if (same) {
            const auto read_opts = [&](const char* key, std::vector<VecAdam>& out) {
                if (!nerf.contains(key) || nerf.at(key).type != Json::Array || nerf.at(key).arr.size() != out.size()) return;
                for (size_t i = 0; i < out.size(); ++i) {
                    const Json& e = nerf.at(key).arr[i]; VecAdam o;
                    o.iter = (uint32_t)e.value("iter", 0.0);
                    if (e.contains("first_moment")) read_vec(e.at("first_moment"), o.m1, 3);
                    if (e.contains("second_moment")) read_vec(e.at("second_moment"), o.m2, 3);
                    if (e.contains("variable")) read_vec(e.at("variable"), o.var, 3);
                    out[i] = o;
                }
            };
            read_opts("cam_pos_offset", m_cam_pos_offset); read_opts("cam_rot_offset", m_cam_rot_offset);
            if (m_n_images) update_transforms(s, 0, m_n_images);
            else for (size_t i = 0; i < m_host_metas.size(); ++i) effective_xform((uint32_t)i, m_host_metas[i].xform, m_host_metas[i].xform_end);
        }
    }
    if (m_mode == MODE_TB_NERF) {
        const Json& g = snap.at("density_grid_binary");
        NGP_CHECK(g.type == Json::Binary, "snapshot: 'density_grid_binary' is not binary");
        const size_t cells = g.str.size() / 2, expect = (size_t)NERF_GRID_N_CELLS * (m_max_cascade + 1);
        if (cells == expect) {
            DeviceBuffer h16;
            binary_to_device(s, g, h16.get<__half>(cells), cells * 2, "density_grid_binary");
            half_to_float(s, cells, h16.as<__half>(), m_density_grid.as<float>());
            // update_density_grid_mean_and_bitfield (testbed_nerf.cu:3538-3567)
            nerf_update_bitfield(s, m_density_grid.as<float>(), m_max_cascade + 1, m_bitfield.as<uint8_t>(), m_mean.as<float>(), m_partials.as<float>());
            nerf_build_coarse_occupancy(s, m_bitfield.as<uint8_t>(), m_coarse.get<uint32_t>(1024));
            NGP_CUDA_CHECK(cudaStreamSynchronize(s));
        } else if (cells != 0) throw Error("Incompatible number of grid cascades.");
    }
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));
}

}  // namespace ngp
