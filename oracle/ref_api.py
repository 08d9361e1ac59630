"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes/numpy front-end of oracle/libngp_ref.so (see oracle/ngp_ref.h).

May be imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

u32, u64, i32, f32, vp = C.c_uint32, C.c_uint64, C.c_int, C.c_float, C.c_void_p


class GridCfg(C.Structure):
    _fields_ = [("n_dims", u32), ("n_levels", u32), ("n_features", u32), ("log2_hashmap_size", u32), ("base_resolution", u32),
                ("per_level_scale", f32)]


class NerfNetCfg(C.Structure):
    _fields_ = [("grid", GridCfg), ("width", u32), ("density_hidden", u32), ("rgb_hidden", u32)]


class AdamCfg(C.Structure):
    _fields_ = [("learning_rate", f32), ("beta1", f32), ("beta2", f32), ("epsilon", f32), ("l2_reg", f32), ("loss_scale", f32),
                ("n_matrix_params", u64), ("optimize_matrix", i32), ("optimize_non_matrix", i32)]


class Image(C.Structure):
    _fields_ = [("pixels", vp), ("data_type", C.c_int32), ("lens_mode", C.c_int32), ("lens_params", f32 * 4), ("res", C.c_int32 * 2),
                ("principal_point", f32 * 2), ("focal_length", f32 * 2), ("xform", f32 * 12), ("xform_end", f32 * 12), ("rolling_shutter", f32 * 4), ("depth", vp)]


class Aabb(C.Structure):
    _fields_ = [("min", f32 * 3), ("max", f32 * 3)]


class SamplerOut(C.Structure):
    _fields_ = [("ray_indices", vp), ("rays", vp), ("numsteps", vp), ("coords", vp), ("ray_counter", u32), ("numsteps_counter", u32)]


class LossCfg(C.Structure):
    _fields_ = [("loss_type", i32), ("rgb_activation", i32), ("density_activation", i32), ("random_bg_color", i32), ("linear_colors", i32),
                ("snap_to_pixel_centers", i32), ("background_color", f32 * 3), ("loss_scale", f32), ("near_distance", f32), ("mean_density", f32), ("color_space", i32), ("exposure", vp), ("exposure_gradient", vp),
                ("depth_supervision_lambda", f32), ("depth_loss_type", i32)]


class LossOut(C.Structure):
    _fields_ = [("dloss_doutput", vp), ("coords_compacted", vp), ("loss", vp), ("numsteps", vp), ("rgb_ray", vp), ("compacted_counter", u32)]


class RenderCfg(C.Structure):
    _fields_ = [("res", C.c_int32 * 2), ("focal_length", f32 * 2), ("screen_center", f32 * 2), ("camera", f32 * 12), ("render_aabb", Aabb),
                ("train_aabb", Aabb), ("cone_angle_constant", f32), ("min_transmittance", f32), ("max_mip", u32), ("spp_index", u32),
                ("snap_to_pixel_centers", i32), ("rgb_activation", i32), ("density_activation", i32), ("linear_colors", i32),
                ("y0", C.c_int32), ("y1", C.c_int32)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "libngp_ref.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libngp_ref.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        l = _LIB
        l.ref_grid_layout.restype = u64
        l.ref_nerf_n_params.restype = u64
        l.ref_nerf_pos_width.restype = u32
        l.ref_mlp_n_params.restype = u32
        l.ref_exp_decay_lr.restype = f32
        l.ref_exp_decay_lr.argtypes = [f32, u32, u32, u32, f32, u32]
        l.ref_update_bitfield.restype = f32
        l.ref_l2_loss.restype = C.c_double
        l.ref_l2_loss.argtypes = [vp, u32, vp, u32, f32, vp]
        l.ref_tcnn_loss.restype = C.c_double
        l.ref_tcnn_loss.argtypes = [i32, vp, u32, vp, u32, u32, f32, vp]
        l.ref_permute.restype = u32; l.ref_permute.argtypes = [u32, u32, u32]
        l.ref_sdf_shuffle.argtypes = [u32, u32, u32, vp, vp, vp, vp]
        l.ref_sdf_normalize.argtypes = [u32, vp, vp, vp, vp, f32, vp, vp]
        l.ref_next_rays_per_batch.restype = u32
        for name in ["ref_to_stepping_space", "ref_from_stepping_space", "ref_calc_dt", "ref_srgb_to_linear", "ref_linear_to_srgb"]:
            getattr(l, name).restype = f32
        l.ref_to_stepping_space.argtypes = [f32, f32]; l.ref_from_stepping_space.argtypes = [f32, f32]; l.ref_calc_dt.argtypes = [f32, f32]
        l.ref_advance_n_steps.restype = f32; l.ref_advance_n_steps.argtypes = [f32, f32, f32]
        l.ref_srgb_to_linear.argtypes = [f32]; l.ref_linear_to_srgb.argtypes = [f32]
        l.ref_ld_random_val.restype = f32; l.ref_ld_random_val.argtypes = [u32, u32]
        l.ref_mip_from_dt.argtypes = [f32, vp, u32]
        l.ref_ema_grid_samples.argtypes = [u32, f32, vp, vp]
        l.ref_generate_grid_samples_nonuniform.argtypes = [u32, u64, u64, u32, vp, vp, vp, vp, u32, f32]
        l.ref_ema_step.argtypes = [u64, f32, u32, vp, vp, vp]
        l.ref_generate_training_samples.argtypes = [u32, vp, u32, u32, u64, u64, u32, vp, vp, u32, i32, f32, vp]
    return _LIB


def P(a):
    """numpy array -> void*"""
    return a.ctypes.data_as(vp) if a is not None else vp(0)


def grid_cfg(n_dims=3, n_levels=16, n_features=2, log2_hashmap_size=19, base_resolution=16, per_level_scale=2.0):
    return GridCfg(n_dims, n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale)


def grid_cfg_from_json(enc, n_dims):
    F = int(enc.get("n_features_per_level", 2))
    L = int(enc.get("n_levels", 16))
    log2T = int(enc.get("log2_hashmap_size", 19))
    base = int(enc.get("base_resolution", 0)) or (1 << (int(enc.get("log2_hashmap_size", 15)) // n_dims))
    return grid_cfg(n_dims, L, F, log2T, base, float(enc.get("per_level_scale", 2.0)))


def nerf_cfg_from_json(cfg):
    return NerfNetCfg(grid_cfg_from_json(cfg["encoding"], 3), 64, int(cfg["network"].get("n_hidden_layers", 1)),
                      int(cfg.get("rgb_network", cfg["network"]).get("n_hidden_layers", 2)))


def grid_layout(cfg):
    L = cfg.n_levels
    off = np.zeros(L + 1, np.uint32); sc = np.zeros(L, np.float32); res = np.zeros(L, np.uint32)
    n = lib().ref_grid_layout(C.byref(cfg), P(off), P(sc), P(res))
    return int(n), off, sc, res


def grid_corners(cfg, x, level):
    n = x.shape[0]; k = 1 << cfg.n_dims
    idx = np.zeros((n, k), np.uint32); w = np.zeros((n, k), np.float32)
    lib().ref_grid_corners(C.byref(cfg), P(x), u32(x.shape[1]), u32(n), u32(level), P(idx), P(w))
    return idx, w


def _with_max_level(setter, max_level, call):
    """run `call` with the per-row level cut-offs installed (GridEncoding::set_max_level_gpu), and clear them again"""
    if max_level is None:
        return call()
    ml = np.ascontiguousarray(max_level, np.float32)
    fn = getattr(lib(), setter); fn.argtypes = [vp]; fn.restype = None
    fn(P(ml))
    try:
        return call()
    finally:
        fn(None)


def grid_forward(cfg, table_h, x, width, max_level=None):
    n = x.shape[0]
    y = np.zeros((n, width), np.float16)
    _with_max_level("ref_grid_set_max_level", max_level,
                    lambda: lib().ref_grid_forward(C.byref(cfg), P(table_h), P(x), u32(x.shape[1]), u32(n), P(y), u32(width), u32(width)))
    return y


def grid_backward(cfg, x, dy, n_params, max_level=None):
    g = np.zeros(n_params, np.float64)
    _with_max_level("ref_grid_set_max_level", max_level,
                    lambda: lib().ref_grid_backward(C.byref(cfg), P(x), u32(x.shape[1]), u32(x.shape[0]), P(dy), u32(dy.shape[1]), P(g)))
    return g


def grid_backward_input(cfg, table_h, x, dy):
    dx = np.zeros((x.shape[0], cfg.n_dims), np.float32)
    lib().ref_grid_backward_input(C.byref(cfg), P(table_h), P(x), u32(x.shape[1]), u32(x.shape[0]), P(dy), u32(dy.shape[1]), P(dx))
    return dx


def mlp_forward(w_h, in_w, n_hidden, x_h, want_acts=False):
    n = x_h.shape[0]
    out = np.zeros((n, 16), np.float16)
    acts = np.zeros((n_hidden, n, 64), np.float16) if want_acts else None
    lib().ref_mlp_forward(P(w_h), u32(in_w), u32(64), u32(n_hidden), u32(16), P(x_h), u32(x_h.shape[1]), u32(n), P(out), u32(16), P(acts))
    return (out, acts) if want_acts else out


def mlp_backward(w_h, in_w, n_hidden, x_h, acts, dout_h):
    n = x_h.shape[0]
    din = np.zeros((n, in_w), np.float16)
    wg = np.zeros(int(lib().ref_mlp_n_params(u32(in_w), u32(64), u32(n_hidden), u32(16))), np.float64)
    lib().ref_mlp_backward(P(w_h), u32(in_w), u32(64), u32(n_hidden), u32(16), P(x_h), u32(x_h.shape[1]), u32(n), P(acts), P(dout_h),
                           u32(dout_h.shape[1]), P(din), u32(in_w), P(wg))
    return din, wg


def nerf_n_params(cfg):
    a, b, c = u64(), u64(), u64()
    n = lib().ref_nerf_n_params(C.byref(cfg), C.byref(a), C.byref(b), C.byref(c))
    return int(n), a.value, b.value, c.value


def nerf_init(cfg, seed=1337):
    p = np.zeros(nerf_n_params(cfg)[0], np.float32)
    lib().ref_nerf_init(C.byref(cfg), u64(seed), P(p))
    return p


def nerf_forward(cfg, params_h, coords, max_level=None):
    out = np.zeros((coords.shape[0], 4), np.float16)
    _with_max_level("ref_nerf_set_max_level", max_level,
                    lambda: lib().ref_nerf_forward(C.byref(cfg), P(params_h), P(coords), u32(coords.shape[0]), P(out)))
    return out


def nerf_density(cfg, params_h, pos):
    out = np.zeros(pos.shape[0], np.float16)
    lib().ref_nerf_density(C.byref(cfg), P(params_h), P(pos), u32(pos.shape[1]), u32(pos.shape[0]), P(out))
    return out


def nerf_backward(cfg, params_h, coords, dL_h, max_level=None):
    g = np.zeros(nerf_n_params(cfg)[0], np.float64)
    _with_max_level("ref_nerf_set_max_level", max_level,
                    lambda: lib().ref_nerf_backward(C.byref(cfg), P(params_h), P(coords), u32(coords.shape[0]), P(dL_h), P(g)))
    return g


def nerf_backward_input(cfg, params_h, coords, dL_h, max_level=None):
    """-> (parameter gradients, dL/d coords [n, 7])"""
    g = np.zeros(nerf_n_params(cfg)[0], np.float64); dc = np.zeros((coords.shape[0], 7), np.float32)
    _with_max_level("ref_nerf_set_max_level", max_level,
                    lambda: lib().ref_nerf_backward_input(C.byref(cfg), P(params_h), P(coords), u32(coords.shape[0]), P(dL_h), P(g), P(dc)))
    return g, dc


def sh4_jacobian(direction):
    J = np.zeros((16, 3), np.float32); d = np.ascontiguousarray(direction, np.float32)
    lib().ref_sh4_jacobian(P(d), P(J))
    return J


def pcg32_floats(seed, skip, n):
    out = np.zeros(n, np.float32)
    lib().ref_pcg32_fill_float(u64(seed), u64(skip), P(out), C.c_size_t(n))
    return out


def adam_step(cfg, master, params_h, grads_h, m, v, steps):
    lib().ref_adam_step(C.byref(cfg), u64(master.size), P(master), P(params_h), P(grads_h), P(m), P(v), P(steps))


def make_image(pixels, data_type, res, focal, xform, principal=(0.5, 0.5), lens_mode=0, lens_params=(0, 0, 0, 0), xform_end=None, rolling_shutter=(0, 0, 0, 0)):
    im = Image()
    im.pixels = pixels.ctypes.data
    im.data_type = data_type; im.lens_mode = lens_mode
    im.lens_params = (f32 * 4)(*lens_params); im.res = (C.c_int32 * 2)(*res)
    im.principal_point = (f32 * 2)(*principal); im.focal_length = (f32 * 2)(*focal)
    im.xform = (f32 * 12)(*[float(v) for v in np.asarray(xform, np.float32).reshape(-1)])
    im.xform_end = (f32 * 12)(*[float(v) for v in np.asarray(xform if xform_end is None else xform_end, np.float32).reshape(-1)])
    im.rolling_shutter = (f32 * 4)(*[float(v) for v in rolling_shutter])
    return im


def make_aabb(lo, hi):
    a = Aabb(); a.min = (f32 * 3)(*lo); a.max = (f32 * 3)(*hi); return a


# ---------------------------------------------------------------------------------------------------------------
# NeRF path wrappers
# ---------------------------------------------------------------------------------------------------------------
def images_from_views(views):
    """views as given to Testbed.set_nerf_dataset -> (ctypes Image array, keepalive list)"""
    arr = (Image * len(views))()
    keep = []
    for i, v in enumerate(views):
        px = v["pixels"]
        px = px.numpy() if hasattr(px, "numpy") else np.ascontiguousarray(px)
        keep.append(px)
        h, w = px.shape[:2]
        arr[i] = make_image(px, 0 if px.dtype == np.uint8 else 1, (w, h), v["focal_length"], v["xform"], v.get("principal_point", (0.5, 0.5)),
                            int(v.get("lens_mode", 0)), v.get("lens_params", (0, 0, 0, 0)), v.get("xform_end"), v.get("rolling_shutter", (0, 0, 0, 0)))
    return arr, keep


def mark_untrained(grid, images):
    lib().ref_mark_untrained_density_grid(u32(grid.size), P(grid), u32(len(images)), images, i32(1))
    return grid


def generate_grid_samples(n, seed, skip, step, aabb, grid_in, n_cascades, thresh):
    pos = np.zeros((n, 3), np.float32); idx = np.zeros(n, np.uint32)
    lib().ref_generate_grid_samples_nonuniform(u32(n), u64(seed), u64(skip), u32(step), C.byref(aabb), P(grid_in), P(pos), P(idx), u32(n_cascades), f32(thresh))
    return pos, idx


def update_bitfield(grid, n_cascades):
    bf = np.zeros(128 ** 3 * 8 // 8, np.uint8)
    mean = lib().ref_update_bitfield(P(grid), u32(n_cascades), P(bf))
    return bf, float(mean)


def _max_level_rand(on, max_level=None, max_level_compacted=None):
    fn = lib().ref_set_max_level_rand; fn.argtypes = [C.c_int, vp, vp]; fn.restype = None
    fn(int(on), None if max_level is None else P(max_level), None if max_level_compacted is None else P(max_level_compacted))


def generate_training_samples(n_rays, aabb, max_samples, seed, skip, images, bitfield, max_mip, snap, cone, max_level_rand=False):
    """max_level_rand (max_level_rand_training): one more draw per ray; the result then carries max_level [max_samples]"""
    out = SamplerOut()
    ray_indices = np.zeros(n_rays, np.uint32); rays = np.zeros((n_rays, 6), np.float32); numsteps = np.zeros((n_rays, 2), np.uint32)
    coords = np.zeros((max_samples, 7), np.float32)
    out.ray_indices, out.rays, out.numsteps, out.coords = P(ray_indices).value, P(rays).value, P(numsteps).value, P(coords).value
    max_level = np.ones(max_samples, np.float32) if max_level_rand else None
    _max_level_rand(max_level_rand, max_level)
    try:
        lib().ref_generate_training_samples(u32(n_rays), C.byref(aabb), u32(max_samples), u32(0), u64(seed), u64(skip), u32(len(images)), images,
                                            P(bitfield), u32(max_mip), i32(int(snap)), f32(cone), C.byref(out))
    finally:
        _max_level_rand(False)
    r = dict(ray_indices=ray_indices, rays=rays, numsteps=numsteps, coords=coords, ray_counter=out.ray_counter, numsteps_counter=out.numsteps_counter)
    if max_level_rand:
        r["max_level"] = max_level
    return r


def compute_loss(n_rays, aabb, seed, skip, max_compacted, rays_counter, cfg, images, net_out_h, ray_indices, rays, numsteps, coords, max_level_rand=False):
    out = LossOut()
    dloss = np.zeros((max_compacted, 4), np.float16); cc = np.zeros((max_compacted, 7), np.float32); loss = np.zeros(n_rays, np.float32)
    ns = numsteps.copy(); rgb_ray = np.zeros((n_rays, 3), np.float32)
    out.dloss_doutput, out.coords_compacted, out.loss, out.numsteps, out.rgb_ray = P(dloss).value, P(cc).value, P(loss).value, P(ns).value, P(rgb_ray).value
    mlc = np.ones(max_compacted, np.float32) if max_level_rand else None
    _max_level_rand(max_level_rand, None, mlc)
    try:
        lib().ref_compute_loss(u32(n_rays), C.byref(aabb), u32(0), u64(seed), u64(skip), u32(max_compacted), u32(rays_counter), C.byref(cfg),
                               u32(len(images)), images, P(net_out_h), P(ray_indices), P(rays), P(coords), C.byref(out))
    finally:
        _max_level_rand(False)
    r = dict(dloss=dloss, coords_compacted=cc, loss=loss, numsteps=ns, rgb_ray=rgb_ray, compacted_counter=out.compacted_counter)
    if max_level_rand:
        r["max_level_compacted"] = mlc
    return r


def set_error_map(cdf_x_cond_y=None, cdf_y=None, cdf_img=None, cdf_res=(0, 0), error_map=None, map_res=(0, 0)):
    """error-map importance sampling state of the sampler / loss / camera-gradient functions; call with no arguments to clear it"""
    fn = lib().ref_set_error_map; fn.argtypes = [vp, vp, vp, C.c_int, C.c_int, vp, C.c_int, C.c_int]; fn.restype = None
    q = lambda a: None if a is None else P(a)
    fn(q(cdf_x_cond_y), q(cdf_y), q(cdf_img), int(cdf_res[0]), int(cdf_res[1]), q(error_map), int(map_res[0]), int(map_res[1]))


def set_distortion(distortion_map=None, gradient=None, gradient_weight=None):
    """distortion map [h, w, 2] (and optional gradient / weight accumulators of the same shape) for the sampler, the tracer and the camera gradient"""
    fn = lib().ref_set_distortion; fn.argtypes = [vp, C.c_int, C.c_int, vp, vp]; fn.restype = None
    q = lambda a: None if a is None else P(a)
    h, w = (0, 0) if distortion_map is None else distortion_map.shape[:2]
    fn(q(distortion_map), int(w), int(h), q(gradient), q(gradient_weight))


def set_cam_gradient_rng(seed, skip, images, snap):
    fn = lib().ref_set_cam_gradient_rng; fn.argtypes = [u64, u64, vp, C.c_int]; fn.restype = None
    fn(seed, skip, C.cast(images, vp) if images is not None else None, int(snap))


def construct_error_cdfs(error_map, n_images, height, width):
    """-> (cdf_x_cond_y [n, h, w], cdf_y [n, h], cdf_img [n]) as Testbed::train_nerf builds them"""
    e = np.ascontiguousarray(error_map, np.float32)
    cxy = np.zeros((n_images, height, width), np.float32); cy = np.zeros((n_images, height), np.float32); cimg = np.zeros(n_images, np.float32)
    fn = lib().ref_construct_error_cdfs; fn.argtypes = [u32, u32, u32, vp, vp, vp, vp]; fn.restype = None
    fn(n_images, height, width, P(e), P(cxy), P(cy), P(cimg))
    return cxy, cy, cimg


def rotmat(v):
    """rotmat(vec3) (common_device.cuh:715-722) -> 3 x 3 (rows)"""
    v = np.ascontiguousarray(v, np.float32); m = np.zeros(9, np.float32)
    fn = lib().ref_rotmat; fn.argtypes = [vp, vp]; fn.restype = None
    fn(P(v), P(m))
    return m.reshape(3, 3).T.copy()


def rotvec(m):
    """rotvec(mat3) (common_device.cuh:698-701), m 3 x 3 (rows)"""
    a = np.ascontiguousarray(np.asarray(m, np.float32).T.reshape(-1)); v = np.zeros(3, np.float32)
    fn = lib().ref_rotvec; fn.argtypes = [vp, vp]; fn.restype = None
    fn(P(a), P(v))
    return v


def compute_cam_gradient(n_rays, aabb, n_images, rays_counter, ray_indices, rays, numsteps, coords, coords_gradient):
    """-> (cam_pos_gradient [n_images, 3], cam_rot_gradient [n_images, 3])"""
    gp = np.zeros((n_images, 3), np.float32); gr = np.zeros((n_images, 3), np.float32)
    fn = lib().ref_compute_cam_gradient
    fn.argtypes = [u32, vp, u32, u32, vp, vp, vp, vp, vp, vp, vp]; fn.restype = None
    fn(n_rays, C.addressof(aabb), n_images, rays_counter, P(ray_indices), P(rays), P(numsteps), P(coords), P(coords_gradient), P(gp), P(gr))
    return gp, gr


def xform_given_rolling_shutter(start, end=None, rolling_shutter=(0, 0, 0, 0), u=0.5, v=0.5, motionblur_time=0.0):
    """get_xform_given_rolling_shutter (common_device.cuh:401-408) -> 12 floats, column major"""
    a = np.ascontiguousarray(np.asarray(start, np.float32).reshape(-1)); b = a if end is None else np.ascontiguousarray(np.asarray(end, np.float32).reshape(-1))
    rs = np.ascontiguousarray(np.asarray(rolling_shutter, np.float32)); out = np.zeros(12, np.float32)
    fn = lib().ref_xform_given_rolling_shutter
    fn.argtypes = [vp, vp, vp, f32, f32, f32, vp]; fn.restype = None
    fn(P(a), P(b), P(rs), float(u), float(v), float(motionblur_time), P(out))
    return out


def render_nerf(rcfg, net_cfg, params_h, bitfield):
    W, H = rcfg.res[0], rcfg.res[1]
    rgba = np.zeros((H, W, 4), np.float32); depth = np.zeros((H, W), np.float32)
    lib().ref_render_nerf(C.byref(rcfg), C.byref(net_cfg), P(params_h), P(bitfield), P(rgba), P(depth))
    return rgba, depth


def synth_views(n_views, w, h, radius, focal, step=1.73205080757 / 1024.0, count=None):
    """CPU-rendered views of the analytic scene (for the CPU reference arm / CPU tests): the first `count` cameras
    of an `n_views`-camera lattice."""
    views = []
    l = lib()
    for i in range(n_views if count is None else count):
        px = np.zeros((h, w), np.uint32)
        l.ref_synth_render_view(u32(i), u32(n_views), f32(radius), f32(focal), i32(w), i32(h), f32(step), P(px))
        xf = np.zeros(12, np.float32)
        l.ref_synth_camera(u32(i), u32(n_views), f32(radius), P(xf))
        views.append({"pixels": px.view(np.uint8).reshape(h, w, 4), "focal_length": (focal, focal), "principal_point": (0.5, 0.5), "xform": xf})
    return views


TCNN_LOSS = {"L2": 0, "RelativeL2": 1, "L1": 2, "MAPE": 3, "SMAPE": 4}


def sdf_shuffle(positions, distances, n_out, seed):
    """first n_out records of tcnn shuffle(seed) over (positions [n,3], distances [n])"""
    positions = np.ascontiguousarray(positions, np.float32); distances = np.ascontiguousarray(distances, np.float32)
    po, do = np.empty((n_out, 3), np.float32), np.empty(n_out, np.float32)
    lib().ref_sdf_shuffle(u32(len(distances)), u32(n_out), u32(seed), P(positions), P(distances), P(po), P(do))
    return po, do


def sdf_normalize(points, distances, raw_min, raw_max, mesh_scale):
    points = np.ascontiguousarray(points, np.float32); distances = np.ascontiguousarray(distances, np.float32)
    lo, hi = np.asarray(raw_min, np.float32), np.asarray(raw_max, np.float32)
    po, do = np.empty_like(points), np.empty_like(distances)
    lib().ref_sdf_normalize(u32(len(distances)), P(points), P(distances), P(lo), P(hi), f32(mesh_scale), P(po), P(do))
    return po, do


def tcnn_loss(kind, pred_h, target, dims, loss_scale=128.0):
    """pred_h fp16 [n, stride]; target fp32 [n, dims] -> (summed loss, dL fp16 [n, stride])"""
    pred_h = np.ascontiguousarray(pred_h, np.float16); target = np.ascontiguousarray(target, np.float32)
    n, stride = pred_h.shape
    dl = np.empty((n, stride), np.float16)
    total = lib().ref_tcnn_loss(i32(TCNN_LOSS[kind]), P(pred_h), u32(stride), P(target), u32(n), u32(dims), f32(loss_scale), P(dl))
    return total, dl
