"""`pyngp`-shaped binding over the C ABI (mirrors src/python_api.cu:266-538 of the reference).

    testbed = Testbed(TestbedMode.Nerf)
    testbed.set_nerf_dataset(views, aabb_scale=1)       # load_training_data with in-memory views (host buffers)
    testbed.reload_network_from_json(config)
    while testbed.frame(): ...                          # one training step per frame, as scripts/run.py:183
    testbed.set_camera_to_training_view(i)              # or testbed.camera_matrix / fov / set_nerf_camera_matrix
    image = testbed.render(w, h, spp, linear=True)

All arithmetic runs in libngp_b200.so; torch only provides the CUDA stream and (for data-parallel training) the
NCCL process group that all-reduces the flat fp16 gradient buffer between train_begin() and train_end().
"""
import ctypes as C
import os
import enum
import json as _json

import numpy as np
import torch

from ._lib import TrainingImage, TrainStats, check, lib


class TestbedMode(enum.IntEnum):      # python_api.cu:271-277 (Volume and None exist as values; this library implements Nerf, Sdf, Image)
    Nerf = 0
    Sdf = 1
    Image = 2
    Volume = 3
    None_ = 4


class GroundTruthRenderMode(enum.IntEnum):   # python_api.cu:282-285
    Shade = 0
    Depth = 1


class RenderMode(enum.IntEnum):       # python_api.cu:287-297
    AO = 0
    Shade = 1
    Normals = 2
    Positions = 3
    Depth = 4
    Distortion = 5
    Cost = 6
    Slice = 7


class ColorSpace(enum.IntEnum):       # python_api.cu:338-341
    Linear = 0
    SRGB = 1


class TonemapCurve(enum.IntEnum):     # python_api.cu:343-348
    Identity = 0
    ACES = 1
    Hable = 2
    Reinhard = 3


class SDFGroundTruthMode(enum.IntEnum):   # python_api.cu:320-324
    RaytracedMesh = 0
    SpheretracedMesh = 1
    SDFBricks = 2


class MeshSdfMode(enum.IntEnum):      # python_api.cu:333-337
    Watertight = 0
    Raystab = 1
    PathEscape = 2


class LossType(enum.IntEnum):         # python_api.cu:311-319
    L2 = 0
    L1 = 1
    Mape = 2
    Smape = 3
    Huber = 4
    SmoothL1 = 4      # legacy name of Huber
    LogL1 = 5
    RelativeL2 = 6


class NerfActivation(enum.IntEnum):   # python_api.cu:321-327
    None_ = 0
    ReLU = 1
    Logistic = 2
    Exponential = 3


class LensMode(enum.IntEnum):         # python_api.cu:350-357 (this library traces Perspective and OpenCV; the loader rejects the others)
    Perspective = 0
    OpenCV = 1
    FTheta = 2
    LatLong = 3
    OpenCVFisheye = 4
    Equirectangular = 5


class BoundingBox:
    """pyngp.BoundingBox (python_api.cu:358-377; include/neural-graphics-primitives/bounding_box.cuh:36-242): axis-aligned box, host side"""
    def __init__(self, a=None, b=None):
        self.min = np.full(3, np.inf, np.float32) if a is None else np.asarray(a, np.float32).copy()
        self.max = np.full(3, -np.inf, np.float32) if b is None else np.asarray(b, np.float32).copy()

    def __repr__(self):
        return f"[min=[{self.min[0]},{self.min[1]},{self.min[2]}], max=[{self.max[0]},{self.max[1]},{self.max[2]}]]"

    def enlarge(self, other):
        if isinstance(other, BoundingBox):
            self.min = np.minimum(self.min, other.min); self.max = np.maximum(self.max, other.max)
        else:
            q = np.asarray(other, np.float32)
            self.min = np.minimum(self.min, q); self.max = np.maximum(self.max, q)

    def inflate(self, amount):
        self.min = self.min - np.float32(amount); self.max = self.max + np.float32(amount)

    def diag(self):
        return self.max - self.min

    def relative_pos(self, pos):
        return (np.asarray(pos, np.float32) - self.min) / self.diag()

    def center(self):
        return np.float32(0.5) * (self.max + self.min)

    def intersection(self, other):
        return BoundingBox(np.maximum(self.min, other.min), np.minimum(self.max, other.max))

    def is_empty(self):
        return bool(np.any(self.max < self.min))

    def intersects(self, other):
        return not self.intersection(other).is_empty()

    def contains(self, p):
        p = np.asarray(p, np.float32)
        return bool(np.all(p >= self.min) and np.all(p <= self.max))

    def distance_sq(self, p):
        p = np.asarray(p, np.float32)
        d = np.maximum(np.maximum(self.min - p, p - self.max), 0.0)
        return float(np.dot(d, d))

    def distance(self, p):
        return float(np.sqrt(self.distance_sq(p)))

    def signed_distance(self, p):
        q = np.abs(np.asarray(p, np.float32) - self.min) - self.diag()
        return float(np.linalg.norm(np.maximum(q, 0.0)) + min(float(q.max()), 0.0))

    def ray_intersect(self, pos, direction):
        """slab test; (tmin, tmax), both FLT_MAX when the ray misses"""
        pos = np.asarray(pos, np.float32); d = np.asarray(direction, np.float32)
        miss = (float(np.finfo(np.float32).max),) * 2
        with np.errstate(divide="ignore", invalid="ignore"):
            t0, t1 = (self.min - pos) / d, (self.max - pos) / d
        lo, hi = np.minimum(t0, t1), np.maximum(t0, t1)
        tmin, tmax = lo[0], hi[0]
        for k in (1, 2):
            if tmin > hi[k] or lo[k] > tmax:
                return miss
            tmin, tmax = max(tmin, lo[k]), min(tmax, hi[k])
        return (float(tmin), float(tmax))

    def get_vertices(self):
        lo, hi = self.min, self.max
        return [np.array([x, y, z], np.float32) for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])]


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_BOOL_OPTS = {"random_bg_color", "linear_colors", "snap_to_pixel_centers", "train_snap_to_pixel_centers", "shall_train_encoding", "shall_train_network",
              "image_snap_to_pixel_centers", "image_linear_colors", "image_stratified", "prefetch_rays", "dp_sharded_optimizer", "device_control_law", "deferred_step_finish",
              "render_ground_truth", "render_with_lens_distortion", "walk_ahead", "optimize_exposure", "max_level_rand_training", "optimize_extrinsics", "optimize_distortion",
              "sample_focal_plane_proportional_to_error", "sample_image_proportional_to_error"}
_FLOAT_OPTS = {"near_distance", "density_grid_decay", "render_min_transmittance", "cone_angle_constant", "exposure", "zoom", "fov", "ground_truth_alpha",
               "render_near_distance", "exposure_l2_reg", "depth_supervision_lambda", "extrinsic_l2_reg", "extrinsic_learning_rate"}
_INT_OPTS = {"loss_type", "rgb_activation", "density_activation", "seed", "sampler_lanes_per_ray", "prefetch_hits", "fov_axis", "walk_ahead_hits", "walk_ahead_misses", "n_steps_between_cam_updates", "n_steps_between_error_map_updates"}
# enum-valued members (python_api.cu:470,522-523,686): reads come back as the enum
_ENUM_OPTS = {}     # filled below the enum definitions
_VECTOR_OPTS = {"screen_center": 2, "relative_focal_length": 2, "up_dir": 3, "background_color": 4}
# members of Testbed itself, which Testbed::set_mode leaves alone (testbed.cu:178-226 resets m_nerf / m_sdf / m_image and the camera): they may be
# set before the mode is known and survive a change of mode
_PERSISTENT_OPTS = {"exposure", "render_mode", "color_space", "tonemap_curve", "snap_to_pixel_centers", "shall_train_encoding", "shall_train_network",
                    "render_ground_truth", "ground_truth_alpha", "background_color", "seed"}


def _default_root_dir():
    """get_root_dir (common.cu:115-133): the first of ".", ".." that holds a "scripts" directory (the reference also looks beside its executable)"""
    for d in (".", ".."):
        if os.path.isdir(os.path.join(d, "scripts")):
            return d
    return ""


def _loads_with_comments(text):
    """json::parse(f, nullptr, true, ignore_comments = true) (testbed.cu:268): // and /* */ comments outside strings are skipped"""
    out, i, n = [], 0, len(text)
    while i < n:
        c = text[i]
        if c == '"':
            j = i + 1
            while j < n and text[j] != '"':
                j += 2 if text[j] == "\\" else 1
            out.append(text[i:j + 1]); i = j + 1
        elif text.startswith("//", i):
            while i < n and text[i] != "\n":
                i += 1
        elif text.startswith("/*", i):
            j = text.find("*/", i + 2)
            i = n if j < 0 else j + 2
        else:
            out.append(c); i += 1
    return _json.loads("".join(out))


def mode_from_scene(path):
    """mode_from_scene (common.cu:144-159)"""
    m = C.c_int32()
    check(lib().ngp_mode_from_scene(os.fspath(path).encode(), C.byref(m)))
    return TestbedMode(m.value)


def builtin_base_config(mode):
    """the hyper-parameters of the reference's configs/<mode>/base.json as the library carries them (used when no config file is found)"""
    t = C.c_char_p()
    check(lib().ngp_builtin_base_config(int(mode), C.byref(t)))
    return _json.loads(t.value.decode())


class RandomMode(enum.IntEnum):       # python_api.cu:299-305
    Random = 0
    Halton = 1
    Sobol = 2
    Stratified = 3


_ENUM_OPTS.update({"depth_loss_type": LossType, "render_mode": RenderMode, "color_space": ColorSpace, "tonemap_curve": TonemapCurve, "image_random_mode": RandomMode})


class _Namespace:
    """One level of the reference binding's attribute tree (`testbed.nerf.training.near_distance`, python_api.cu:550-695): attribute
    reads and writes are forwarded to the flat option of the owning Testbed named in `fields`; `extra` holds child namespaces and
    bound methods."""
    def __init__(self, tb, fields, extra=None):
        object.__setattr__(self, "_tb", tb)
        object.__setattr__(self, "_fields", fields)
        object.__setattr__(self, "_extra", extra or {})

    def __getattr__(self, name):
        if name in self._extra:
            v = self._extra[name]
            return v() if isinstance(v, _Lazy) else v
        if name in self._fields:
            return getattr(self._tb, self._fields[name])
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name not in self._fields:
            raise AttributeError(f"{name} is not a writable member on this path")
        setattr(self._tb, self._fields[name], value)


class _Lazy:
    def __init__(self, fn):
        self.fn = fn

    def __call__(self):
        return self.fn()


class _Dataset:
    """nerf.training.dataset (NerfDataset, python_api.cu:594-610): n_images, metadata, transforms, scale, offset, aabb_scale, up, render_aabb, paths"""
    def __init__(self, tb):
        self._tb = tb

    @property
    def n_images(self):
        return self._tb.n_training_views()

    @property
    def metadata(self):
        tb = self._tb

        class M:
            def __init__(s, i, v):
                s.resolution = list(v["res"]); s.focal_length = list(v["focal_length"]); s.principal_point = list(v["principal_point"])
                rs = (C.c_float * 4)()
                check(lib().ngp_testbed_training_view_motion(tb._h, i, None, None, rs))
                s.rolling_shutter = list(rs)
        return [M(i, self._tb.training_view(i)) for i in range(self.n_images)]

    @property
    def transforms(self):
        """TrainingXForm per view: .start / .end, the poses as given (3 x 4, NGP convention)"""
        class X:
            pass
        out = []
        for i in range(self.n_images):
            a, b = (C.c_float * 12)(), (C.c_float * 12)()
            check(lib().ngp_testbed_training_view_motion(self._tb._h, i, a, b, None))
            x = X(); x.start = np.array(list(a), np.float32).reshape(4, 3).T.copy(); x.end = np.array(list(b), np.float32).reshape(4, 3).T.copy()
            out.append(x)
        return out

    @property
    def scale(self):
        return float(self._tb.get_option("dataset_scale"))

    @property
    def aabb_scale(self):
        return int(self._tb.get_option("dataset_aabb_scale"))

    @property
    def offset(self):
        return self._tb._get_vec("dataset_offset", 3)

    @property
    def up(self):
        return self._tb.up_dir

    @property
    def render_aabb(self):
        return self._tb.render_aabb

    @property
    def paths(self):
        return [""] * self.n_images            # the pixels arrive decoded (ngp_nerf_dataset_set_image): the library keeps no file names


class Testbed:
    def __init__(self, mode=TestbedMode.None_, data_path=None, network_config=None):
        """Testbed() / Testbed(mode) / Testbed(mode, data_path, config_path | json) (python_api.cu:388-390).  Without a mode the testbed is
        an empty shell, as in the reference (ETestbedMode::None): load_training_data / load_snapshot / load_file pick the mode from what
        they are given (testbed.cu:144-150,4950-4955) and only then does the library-side testbed exist."""
        object.__setattr__(self, "_handle", C.c_void_p())
        object.__setattr__(self, "mode", TestbedMode.None_)
        object.__setattr__(self, "_explicit", {})                   # every option the user has set: replayed onto the testbed of the next mode
        object.__setattr__(self, "_root_dir", _default_root_dir())
        object.__setattr__(self, "_requested_config", None)         # reload_network_from_file before the mode is known (testbed.cu:288-292)
        object.__setattr__(self, "training_batch_size", 1 << 18)    # testbed.h:1005
        object.__setattr__(self, "shall_train", True)
        object.__setattr__(self, "_keep", [])
        object.__setattr__(self, "_dist", None)
        training = _Namespace(self, {"random_bg_color": "random_bg_color", "loss_type": "loss_type", "near_distance": "near_distance",
                                     "density_grid_decay": "density_grid_decay", "linear_colors": "linear_colors",
                                     "snap_to_pixel_centers": "train_snap_to_pixel_centers", "view": "training_view_index",
                                     "optimize_exposure": "optimize_exposure", "exposure_l2_reg": "exposure_l2_reg",
                                     "n_steps_between_cam_updates": "n_steps_between_cam_updates",
                                     "depth_supervision_lambda": "depth_supervision_lambda", "depth_loss_type": "depth_loss_type",
                                     "optimize_extrinsics": "optimize_extrinsics", "optimize_distortion": "optimize_distortion", "extrinsic_l2_reg": "extrinsic_l2_reg",
                                     "extrinsic_learning_rate": "extrinsic_learning_rate",
                                     "sample_focal_plane_proportional_to_error": "sample_focal_plane_proportional_to_error",
                                     "sample_image_proportional_to_error": "sample_image_proportional_to_error",
                                     "n_steps_between_error_map_updates": "n_steps_between_error_map_updates"},
                              {"set_image": self.set_training_image, "set_camera_intrinsics": self.set_camera_intrinsics,
                               "set_camera_extrinsics": self.set_camera_extrinsics, "get_camera_extrinsics": self.get_camera_extrinsics,
                               "set_camera_extrinsics_rolling_shutter": self.set_camera_extrinsics_rolling_shutter, "dataset": _Dataset(self),
                               "n_images_for_training": _Lazy(self.n_training_views)})
        object.__setattr__(self, "nerf", _Namespace(self, {"rgb_activation": "rgb_activation", "density_activation": "density_activation",
                                                           "cone_angle_constant": "cone_angle_constant",
                                                           "render_min_transmittance": "render_min_transmittance",
                                                           "render_with_lens_distortion": "render_with_lens_distortion", "sharpen": "nerf_sharpen"},
                                                    {"training": training}))
        image_training = _Namespace(self, {"snap_to_pixel_centers": "image_snap_to_pixel_centers", "linear_colors": "image_linear_colors"})
        object.__setattr__(self, "image", _Namespace(self, {"random_mode": "image_random_mode"}, {"training": image_training}))
        self._set_mode(TestbedMode(int(mode)))
        if data_path is not None:
            self.load_training_data(data_path)
        if network_config is not None:
            if isinstance(network_config, (dict,)) or (isinstance(network_config, str) and network_config.lstrip().startswith("{")):
                self.reload_network_from_json(network_config)
            else:
                self.reload_network_from_file(network_config)

    @property
    def _h(self):
        h = self._handle
        if not h:
            raise RuntimeError("This testbed has no mode yet: load_training_data / load_snapshot / load_file decide it, or construct Testbed(mode)")
        return h

    def _set_mode(self, mode):
        """Testbed::set_mode (testbed.cu:178-226): everything mode-specific (data, network, trainer) is dropped, the options stay."""
        mode = TestbedMode(int(mode))
        if mode == self.mode:
            return
        if mode == TestbedMode.Volume:
            raise RuntimeError("Volume mode (NanoVDB scenes) is not part of this library")
        if self._handle:
            lib().ngp_testbed_destroy(self._handle)
            object.__setattr__(self, "_handle", C.c_void_p())
        object.__setattr__(self, "mode", TestbedMode.None_)
        object.__setattr__(self, "_dist", None)
        if mode == TestbedMode.None_:
            return
        h = C.c_void_p()
        check(lib().ngp_testbed_create(int(mode), C.byref(h)))
        object.__setattr__(self, "_handle", h)
        object.__setattr__(self, "mode", mode)
        check(lib().ngp_testbed_set_root_dir(h, os.fspath(self._root_dir).encode()))
        for name, value in list(self._explicit.items()):
            setattr(self, name, value)
        if self._requested_config is not None:
            req = self._requested_config
            object.__setattr__(self, "_requested_config", None)
            self.reload_network_from_file(req)

    def __del__(self):
        h = self.__dict__.get("_handle", None)
        if h and lib is not None:     # `lib` is already torn down when the interpreter exits
            try:
                lib().ngp_testbed_destroy(h)
            except Exception:
                pass
            object.__setattr__(self, "_handle", None)

    # ---- runtime knobs (def_readwrite members, python_api.cu:452-537) ----
    def _get_vec(self, name, n):
        buf = (C.c_float * n)()
        check(lib().ngp_testbed_get_vector(self._h, name.encode(), buf, n))
        return np.array(list(buf), np.float32)

    def _set_vec(self, name, values, n):
        v = np.asarray(values, np.float32).reshape(-1)
        if v.size != n:
            raise ValueError(f"{name} takes {n} components")
        check(lib().ngp_testbed_set_vector(self._h, name.encode(), (C.c_float * n)(*[float(x) for x in v]), n))

    def __getattr__(self, name):
        if name == "root_dir":               # python_api.cu:534-537
            return self._root_dir
        if name == "nerf_sharpen":           # nerf.sharpen (python_api.cu:556): the training images are used as they are
            return float(self.__dict__.get("_explicit", {}).get(name, 0.0))
        if name.startswith("_"):
            raise AttributeError(name)
        if not self.__dict__.get("_handle") and name in self.__dict__.get("_explicit", {}):
            return self._explicit[name]
        if name in _VECTOR_OPTS:
            return self._get_vec(name, _VECTOR_OPTS[name])
        if name in ("aabb", "render_aabb"):
            v = self._get_vec(name, 6)
            return BoundingBox(v[:3], v[3:])
        if name == "camera_matrix":          # m_camera as pybind hands it out: a 3 x 4 array (rows), [R | t]
            return self._get_vec("camera_matrix", 12).reshape(4, 3).T.copy()
        if name in ("render_groundtruth",):
            return self.render_ground_truth
        if name == "scale":
            return self.get_option("scale")
        if name == "training_view_index":    # nerf.training.view (python_api.cu:640): the view set_camera_to_training_view selected
            return int(self.get_option("training_view"))
        if name == "fov_xy":                 # focal_length_to_fov per axis (testbed.cu:3818-3820)
            f = self._get_vec("relative_focal_length", 2)
            return np.degrees(2.0 * np.arctan(1.0 / (2.0 * f))).astype(np.float32)
        if name == "view_dir":
            return self.camera_matrix[:, 2].copy()
        if name == "view_pos":
            return self.camera_matrix[:, 3].copy()
        if name == "look_at":                # testbed.cu:793-795
            m = self.camera_matrix
            return m[:, 3] + m[:, 2] * np.float32(self.scale)
        if name in _ENUM_OPTS:
            return _ENUM_OPTS[name](int(self.get_option(name)))
        if name in _BOOL_OPTS | _FLOAT_OPTS | _INT_OPTS:
            v = C.c_double()
            check(lib().ngp_testbed_get_option(self._h, name.encode(), C.byref(v)))
            return bool(v.value) if name in _BOOL_OPTS else (float(v.value) if name in _FLOAT_OPTS else int(v.value))
        raise AttributeError(name)

    def get_option(self, name):
        """any option of ngp_testbed_get_option by name (incl. read-only diagnostics such as "l2_window_bytes")"""
        v = C.c_double()
        check(lib().ngp_testbed_get_option(self._h, name.encode(), C.byref(v)))
        return v.value

    def __setattr__(self, name, value):
        if name == "root_dir":
            object.__setattr__(self, "_root_dir", os.fspath(value))
            if self._handle:
                check(lib().ngp_testbed_set_root_dir(self._handle, os.fspath(value).encode()))
            return
        if name == "nerf_sharpen":
            if float(value) != 0.0:
                raise RuntimeError("nerf.sharpen: sharpening of the training images is not part of this library (leave it at 0)")
            self._explicit[name] = 0.0
            return
        if name in _PERSISTENT_OPTS:
            self._explicit[name] = value
            if not self._handle:             # no mode yet: kept, and applied when a mode is chosen
                return
        if name in _BOOL_OPTS | _FLOAT_OPTS | _INT_OPTS | set(_ENUM_OPTS):
            check(lib().ngp_testbed_set_option(self._h, name.encode(), float(value)))
        elif name == "background_color":     # vec4 in the reference; three components keep the alpha that is set
            v = list(np.asarray(value, np.float32).reshape(-1))
            self._set_vec(name, v if len(v) == 4 else v + [float(self.background_color[3])], 4)
        elif name in _VECTOR_OPTS:
            self._set_vec(name, value, _VECTOR_OPTS[name])
        elif name in ("aabb", "render_aabb"):
            self._set_vec(name, np.concatenate([np.asarray(value.min, np.float32), np.asarray(value.max, np.float32)]), 6)
        elif name == "camera_matrix":        # 3 x 4 (rows) or 4 x 4, NGP convention
            m = np.asarray(value, np.float32)
            if m.shape not in ((3, 4), (4, 4)):
                raise ValueError("camera_matrix must be a 3x4 or 4x4 matrix")
            self._set_vec(name, np.ascontiguousarray(m[:3, :4].T).reshape(-1), 12)
        elif name == "render_groundtruth":
            self.render_ground_truth = value
        elif name == "training_view_index":
            check(lib().ngp_testbed_set_option(self._h, b"training_view", float(int(value))))
        elif name == "fov_xy":               # fov_to_focal_length per axis (testbed.cu:3822-3824)
            self._set_vec("relative_focal_length", 0.5 / np.tan(0.5 * np.radians(np.asarray(value, np.float64))), 2)
        elif name == "scale":                # set_scale (testbed.cu:801-805): the eye moves on the view axis, look_at stays
            m, old = self.camera_matrix, self.scale
            look = m[:, 3] + m[:, 2] * np.float32(old)
            m[:, 3] = (m[:, 3] - look) * np.float32(float(value) / old) + look
            self.camera_matrix = m
            check(lib().ngp_testbed_set_option(self._h, b"scale", float(value)))
        elif name == "look_at":              # set_look_at (testbed.cu:797-799)
            m = self.camera_matrix
            m[:, 3] += np.asarray(value, np.float32) - self.look_at
            self.camera_matrix = m
        elif name == "view_dir":             # set_view_dir (testbed.cu:807-813)
            d = np.asarray(value, np.float32); look = self.look_at; m = self.camera_matrix
            unit = lambda v: v / np.linalg.norm(v)
            m[:, 0] = unit(np.cross(d, self.up_dir)); m[:, 1] = unit(np.cross(d, m[:, 0])); m[:, 2] = unit(d)
            self.camera_matrix = m
            self.look_at = look
        else:
            object.__setattr__(self, name, value)

    # ---- view state (python_api.cu:483-491) ----
    def reset_camera(self):
        check(lib().ngp_testbed_reset_camera(self._h))

    def set_nerf_camera_matrix(self, cam):
        """camera-to-world of the original NeRF convention, 3 x 4 (rows) or 4 x 4; converted with the data set's scale / offset"""
        m = np.asarray(cam, np.float32)
        if m.shape not in ((3, 4), (4, 4)):
            raise ValueError("cam must be a 3x4 or 4x4 matrix")
        col = np.ascontiguousarray(m[:3, :4].T).reshape(-1)
        check(lib().ngp_testbed_set_nerf_camera_matrix(self._h, col.ctypes.data_as(C.c_void_p)))

    def set_camera_to_training_view(self, trainview):
        check(lib().ngp_testbed_set_camera_to_training_view(self._h, int(trainview)))

    def first_training_view(self):
        self.set_camera_to_training_view(0)

    def last_training_view(self):
        self.set_camera_to_training_view(self.n_training_views() - 1)

    def previous_training_view(self):
        self.set_camera_to_training_view(max(self.training_view_index - 1, 0))

    def next_training_view(self):
        self.set_camera_to_training_view(min(self.training_view_index + 1, self.n_training_views() - 1))

    def reset_accumulation(self, due_to_camera_movement=False, immediate_redraw=True):
        """every render() call starts from an empty accumulator (render_to_cpu resets it as well, python_api.cu:129): nothing to do"""

    # ---- data ----
    def set_nerf_dataset(self, views, aabb_scale=1):
        """views: list of dicts {pixels: np.uint8 [H,W,4] | np.float32 [H,W,4] (host), focal_length: (fx,fy) px,
        xform: 12 floats (NGP mat4x3 column major), principal_point=(.5,.5), lens_mode, lens_params}."""
        arr = (TrainingImage * len(views))()
        for i, v in enumerate(views):
            px = v["pixels"]
            if isinstance(px, torch.Tensor):
                assert not px.is_cuda, "views are host buffers (pinned torch tensors or numpy arrays)"
                dtype_byte = px.dtype == torch.uint8
                h, w = px.shape[:2]
                ptr = px.data_ptr()
            else:
                px = np.ascontiguousarray(px)
                dtype_byte = px.dtype == np.uint8
                h, w = px.shape[:2]
                ptr = px.ctypes.data
            self._keep.append(px)
            arr[i].pixels = ptr
            arr[i].data_type = 0 if dtype_byte else 1
            arr[i].res = (C.c_int32 * 2)(w, h)
            arr[i].focal_length = (C.c_float * 2)(*v["focal_length"])
            arr[i].principal_point = (C.c_float * 2)(*v.get("principal_point", (0.5, 0.5)))
            arr[i].lens_mode = int(v.get("lens_mode", 0))
            arr[i].lens_params = (C.c_float * 4)(*v.get("lens_params", (0, 0, 0, 0)))
            arr[i].xform = (C.c_float * 12)(*[float(x) for x in np.asarray(v["xform"], np.float32).reshape(-1)])
        check(lib().ngp_testbed_set_nerf_dataset(self._h, _stream(), len(views), arr, int(aabb_scale)))
        self._keep.clear()
        for i, v in enumerate(views):      # optional: "xform_end" (NGP convention, like "xform") and "rolling_shutter" (A, B, C, D)
            if v.get("xform_end") is not None or any(float(x) != 0.0 for x in v.get("rolling_shutter", (0, 0, 0, 0))):
                to34 = lambda m: np.asarray(m, np.float32).reshape(4, 3).T
                self.set_camera_extrinsics_rolling_shutter(i, to34(v["xform"]), to34(v["xform"] if v.get("xform_end") is None else v["xform_end"]),
                                                           v.get("rolling_shutter", (0, 0, 0, 0)), convert_to_ngp=False)

    def load_training_data(self, path):
        """Testbed.load_training_data(path) (python_api.cu:398, testbed.cu:135-165).  The mode follows the scene (mode_from_scene): a
        transforms.json, a directory holding *.json files (or, an extension of this binding, a list of either) is a NeRF scene; any other
        file is an image to learn in Image mode.  The NeRF loader logic (frame order, culling, intrinsics, NeRF -> NGP poses) runs in the
        library (csrc/nerf_dataset.cu); this binding only decodes the image files it names (OpenCV: png / jpg / exr ...)."""
        first = path if isinstance(path, (str, os.PathLike)) else list(path)[0]
        scene_mode = mode_from_scene(first)
        if scene_mode == TestbedMode.None_:
            raise RuntimeError(f"Data path '{os.fspath(first)}' does not exist.")
        if scene_mode == TestbedMode.Volume:
            raise RuntimeError("Volume mode (NanoVDB scenes) is not part of this library")
        self._set_mode(scene_mode)
        if scene_mode == TestbedMode.Image:
            return self._load_image(os.fspath(first))
        if scene_mode == TestbedMode.Sdf:
            raise RuntimeError("load_mesh: mesh files are not read by this library; hand the (point, distance) records to override_sdf_training_data")
        os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")   # OpenCV reads *.exr only when this is set before its first import
        import cv2
        paths = []
        for p in ([path] if isinstance(path, (str, os.PathLike)) else list(path)):
            p = os.fspath(p)
            if os.path.isdir(p):
                paths += sorted(os.path.join(p, f) for f in os.listdir(p) if f.lower().endswith(".json"))
            else:
                paths.append(p)
        d = C.c_void_p()
        check(lib().ngp_nerf_dataset_open("\n".join(paths).encode(), C.byref(d)))
        try:
            buf = C.create_string_buffer(4096)
            for i in range(lib().ngp_nerf_dataset_n_images(d)):
                check(lib().ngp_nerf_dataset_image_path(d, i, buf, 4096))
                f = buf.value.decode()
                img = cv2.imread(f, cv2.IMREAD_UNCHANGED)
                if img is None:
                    raise RuntimeError(f"Could not find image file '{f}'.")
                if img.ndim == 2:
                    img = np.repeat(img[..., None], 3, axis=2)
                is_float = img.dtype == np.float32
                if img.dtype == np.uint16:
                    img = (img >> 8).astype(np.uint8)
                if img.shape[2] == 3:      # stb_image with 4 requested components: opaque alpha
                    alpha = np.full(img.shape[:2] + (1,), 1.0 if is_float else 255, img.dtype)
                    img = np.concatenate([img, alpha], axis=2)
                img = np.ascontiguousarray(img[..., [2, 1, 0, 3]])          # BGRA -> RGBA
                check(lib().ngp_nerf_dataset_set_image(d, i, img.ctypes.data_as(C.c_void_p), int(is_float), img.shape[1], img.shape[0]))
                check(lib().ngp_nerf_dataset_depth_path(d, i, buf, 4096))
                if buf.value:                   # "integer_depth_scale" + "depth_path" (nerf_loader.cu:609-621): 16-bit, one channel
                    dimg = cv2.imread(buf.value.decode(), cv2.IMREAD_UNCHANGED)
                    if dimg is None:
                        raise RuntimeError(f"Could not load depth image '{buf.value.decode()}'.")
                    if dimg.ndim == 3:
                        dimg = dimg[..., 0]
                    if dimg.dtype == np.uint8:  # stbi_load_16 widens 8-bit files
                        dimg = dimg.astype(np.uint16) * 257
                    dimg = np.ascontiguousarray(dimg, np.uint16)
                    check(lib().ngp_nerf_dataset_set_depth(d, i, dimg.ctypes.data_as(C.c_void_p), dimg.shape[1], dimg.shape[0]))
            check(lib().ngp_testbed_load_nerf_dataset(self._h, _stream(), d))
        finally:
            lib().ngp_nerf_dataset_close(d)

    def _load_image(self, path):
        """Testbed::load_image (testbed_image.cu:372-437): *.exr and other float formats are kept as they are (linear RGBA), *.bin is the
        reference's raw dump (int32 height, int32 width, then RGBA halves), everything else is an 8-bit file whose sRGB bytes the library
        turns into linear premultiplied floats on the device, as load_stbi_gpu does."""
        if path.lower().endswith(".bin"):
            with open(path, "rb") as f:
                h, w = np.frombuffer(f.read(8), np.int32)
                px = np.frombuffer(f.read(int(h) * int(w) * 8), np.float16).reshape(int(h), int(w), 4).astype(np.float32)
            return self.set_image(px)
        os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")   # OpenCV reads *.exr only when this is set before its first import
        import cv2
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        if img is None:
            raise RuntimeError(f"Image file '{path}' could not be read.")
        if img.ndim == 2:
            img = np.repeat(img[..., None], 3, axis=2)
        if img.dtype == np.uint16:
            img = (img >> 8).astype(np.uint8)
        is_float = img.dtype != np.uint8
        if img.shape[2] == 3:
            img = np.concatenate([img, np.full(img.shape[:2] + (1,), 1.0 if is_float else 255, img.dtype)], axis=2)
        img = np.ascontiguousarray(img[..., [2, 1, 0, 3]])              # BGRA -> RGBA
        if is_float:
            return self.set_image(img.astype(np.float32))
        check(lib().ngp_testbed_set_image_rgba8(self._h, _stream(), img.ctypes.data_as(C.c_void_p), img.shape[1], img.shape[0]))

    def set_training_image(self, frame_idx, pixels, depth_img=None, depth_scale=1.0):
        """nerf.training.set_image(frame_idx, img, depth_img, depth_scale) (python_api.cu:49-67): host pixels -> the device copy of one view; with
        depth_img ([h, w] float32, or uint16) also the view's depth map, kept as value * depth_scale (scene units)."""
        if isinstance(pixels, torch.Tensor):
            assert not pixels.is_cuda and pixels.is_contiguous()
            ptr = pixels.data_ptr()
        else:
            pixels = np.ascontiguousarray(pixels); ptr = pixels.ctypes.data
        check(lib().ngp_testbed_set_training_image(self._h, _stream(), int(frame_idx), C.c_void_p(ptr)))
        if depth_img is not None:
            self.set_training_depth(frame_idx, depth_img, depth_scale)

    def set_training_depth(self, frame_idx, depth_img, depth_scale=1.0):
        """the depth half of NerfDataset::set_training_image (nerf_loader.cu:943-959); depth_img None with depth_scale < 0 drops the view's map"""
        if depth_img is None:
            check(lib().ngp_testbed_set_training_depth(self._h, _stream(), int(frame_idx), None, 1, float(depth_scale)))
            return
        d = np.asarray(depth_img)
        d = np.ascontiguousarray(d, np.uint16 if d.dtype == np.uint16 else np.float32)
        check(lib().ngp_testbed_set_training_depth(self._h, _stream(), int(frame_idx), d.ctypes.data_as(C.c_void_p), 0 if d.dtype == np.uint16 else 1,
                                                   float(depth_scale)))

    def set_camera_intrinsics(self, frame_idx, fx=0.0, fy=0.0, cx=-0.5, cy=-0.5, k1=0.0, k2=0.0, p1=0.0, p2=0.0):
        """nerf.training.set_camera_intrinsics (python_api.cu:583-593)"""
        check(lib().ngp_testbed_set_camera_intrinsics(self._h, _stream(), int(frame_idx), fx, fy, cx, cy, k1, k2, p1, p2))

    def set_camera_extrinsics(self, frame_idx, camera_to_world, convert_to_ngp=True):
        """nerf.training.set_camera_extrinsics (python_api.cu:594-605); camera_to_world: 3x4 (rows) or 4x4 row-major matrix"""
        m = np.asarray(camera_to_world, np.float32)
        if m.shape not in ((3, 4), (4, 4)):
            raise RuntimeError("camera_to_world must be a 3x4 or 4x4 matrix")
        col_major = np.ascontiguousarray(m[:3, :4].T).reshape(-1)
        check(lib().ngp_testbed_set_camera_extrinsics(self._h, _stream(), int(frame_idx), col_major.ctypes.data_as(C.c_void_p), int(convert_to_ngp)))

    def get_camera_extrinsics(self, frame_idx, convert_to_ngp=True):
        """nerf.training.get_camera_extrinsics(frame_idx, convert_to_ngp=True) (python_api.cu:594-603): the view's pose as trained (3 x 4, rows),
        in the NeRF convention of transforms.json unless convert_to_ngp is False"""
        out = (C.c_float * 12)()
        check(lib().ngp_testbed_get_camera_extrinsics(self._h, int(frame_idx), int(convert_to_ngp), out))
        return np.array(list(out), np.float32).reshape(4, 3).T.copy()

    def set_camera_extrinsics_rolling_shutter(self, frame_idx, camera_to_world_start, camera_to_world_end, rolling_shutter, convert_to_ngp=True):
        """nerf.training.set_camera_extrinsics_rolling_shutter (python_api.cu:606-617): 3x4 (rows) or 4x4 matrices, rolling_shutter = (A, B, C, D)"""
        mats = []
        for m in (camera_to_world_start, camera_to_world_end):
            m = np.asarray(m, np.float32)
            if m.shape not in ((3, 4), (4, 4)):
                raise RuntimeError("camera_to_world must be a 3x4 or 4x4 matrix")
            mats.append(np.ascontiguousarray(m[:3, :4].T).reshape(-1))
        rs = np.ascontiguousarray(np.asarray(rolling_shutter, np.float32).reshape(-1))
        if rs.size != 4:
            raise RuntimeError("rolling_shutter must have 4 components")
        check(lib().ngp_testbed_set_camera_extrinsics_rolling_shutter(self._h, _stream(), int(frame_idx), mats[0].ctypes.data_as(C.c_void_p),
                                                                      mats[1].ctypes.data_as(C.c_void_p), rs.ctypes.data_as(C.c_void_p), int(convert_to_ngp)))

    def reset(self, reset_density_grid=True):
        check(lib().ngp_testbed_reset(self._h, int(reset_density_grid)))

    def set_image(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.float32)
        assert rgba.ndim == 3 and rgba.shape[2] == 4
        check(lib().ngp_testbed_set_image(self._h, _stream(), rgba.ctypes.data_as(C.c_void_p), rgba.shape[1], rgba.shape[0]))

    def override_sdf_training_data(self, points, distances, raw_aabb=((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)), mesh_scale=1.0):
        """Testbed.override_sdf_training_data(points [n,3], distances [n]) (python_api.cu:69-99).  raw_aabb / mesh_scale are what
        the reference takes from the loaded mesh; the defaults mean "already in the unit cube"."""
        points = np.ascontiguousarray(points, np.float32); distances = np.ascontiguousarray(distances, np.float32)
        if points.ndim != 2 or distances.ndim != 1 or points.shape[0] != distances.shape[0] or points.shape[1] != 3:
            raise RuntimeError("Invalid Points<->Distances data")
        lo, hi = (C.c_float * 3)(*raw_aabb[0]), (C.c_float * 3)(*raw_aabb[1])
        check(lib().ngp_testbed_override_sdf_training_data(self._h, _stream(), points.ctypes.data_as(C.c_void_p), distances.ctypes.data_as(C.c_void_p),
                                                           points.shape[0], lo, hi, float(mesh_scale)))

    # ---- network ----
    def reload_network_from_json(self, config, config_base_path=""):
        """python_api.cu:402-407; a config with a "parent" key is merged over <directory of config_base_path>/<parent> (testbed.cu:95-106)"""
        cfg = config if isinstance(config, str) else _json.dumps(config)
        check(lib().ngp_testbed_reload_network_from_json_base(self._h, cfg.encode(), os.fspath(config_base_path).encode()))

    def reload_network_from_file(self, path=""):
        """python_api.cu:400-401, testbed.cu:275-309: `path`, else <root_dir>/configs/<mode>/<path>; "" reloads the last one ("base.json" at first).
        Before the mode is known the request is only remembered (:288-292)."""
        if not self._handle:
            if path:
                object.__setattr__(self, "_requested_config", os.fspath(path))
            return
        check(lib().ngp_testbed_reload_network_from_file(self._h, os.fspath(path).encode()))

    def n_params(self):
        return lib().ngp_testbed_n_params(self._h)

    def n_encoding_params(self):
        return lib().ngp_testbed_n_encoding_params(self._h)

    # ---- snapshots (python_api.cu:446-449) ----
    def sync_inference_params(self):
        """COLLECTIVE (sharded optimizer only, otherwise a no-op): fetch the other ranks' slices of the EMA / inference parameters"""
        check(lib().ngp_testbed_sync_inference_params(self._h, _stream()))

    def save_snapshot(self, path, include_optimizer_state=False, compress=True):
        self.sync_inference_params()
        check(lib().ngp_testbed_save_snapshot(self._h, _stream(), os.fspath(path).encode(), int(include_optimizer_state), int(compress)))

    def load_snapshot(self, path):
        """Testbed::load_snapshot (testbed.cu:4931-5057): the testbed switches to the mode the snapshot was saved in (:4950-4955)"""
        p = os.fspath(path)
        m = C.c_int32()
        check(lib().ngp_snapshot_mode(p.encode(), C.byref(m)))
        self._set_mode(TestbedMode(m.value))
        check(lib().ngp_testbed_load_snapshot(self._h, _stream(), p.encode()))

    def load_file(self, path):
        """Testbed::load_file (testbed.cu:316-375): snapshots by extension; a *.json is looked into (snapshot / network config / camera
        path); anything else is training data."""
        p = os.fspath(path)
        if not os.path.exists(p):
            if p.lower().endswith(".json") and self._handle:          # a name that resolves under <root_dir>/configs/<mode>/
                return self.reload_network_from_file(p)
            raise RuntimeError(f"File '{p}' does not exist.")
        if p.lower().endswith((".ingp", ".msgpack")):
            return self.load_snapshot(p)
        if p.lower().endswith(".json"):
            with open(p) as f:
                doc = _loads_with_comments(f.read())
            if "snapshot" in doc:
                raise RuntimeError("load_file: snapshots in JSON form are not supported; use .ingp / .msgpack")
            if any(k in doc for k in ("parent", "network", "encoding", "loss", "optimizer")):
                return self.reload_network_from_file(p)
            if "path" in doc:
                return self.load_camera_path(p)
        self.load_training_data(p)
        self.shall_train = True

    # ---- what the reference binding has and a headless compute library does not (python_api.cu:408-414,428-445) ----
    def want_repl(self):
        return False                         # no window, no key to ask for a REPL with (python_api.cu:413)

    def init_window(self, *args, **kwargs):
        raise RuntimeError("init_window: the GUI (GLFW / OpenGL) is unavailable on this box; use frame() / train() and render()")

    def init_vr(self):
        raise RuntimeError("init_vr: OpenXR is unavailable on this box")

    def load_camera_path(self, path):
        raise RuntimeError("load_camera_path: camera paths (src/camera_path.cu) are not part of this library; "
                           "set camera_matrix / set_nerf_camera_matrix per frame and call render()")

    def compute_and_save_marching_cubes_mesh(self, *args, **kwargs):
        raise RuntimeError("compute_and_save_marching_cubes_mesh: marching cubes (src/marching_cubes.cu) is not part of this library")

    def compute_marching_cubes_mesh(self, *args, **kwargs):
        raise RuntimeError("compute_marching_cubes_mesh: marching cubes (src/marching_cubes.cu) is not part of this library")

    # ---- training ----
    def enable_data_parallel(self, process_group=None, native=None):
        """Shard training rays over the ranks of a torch.distributed process group (collective: every rank calls it).
        native (default with an NCCL group): the library gets its own NCCL communicator (the unique id travels over the
        process group) and train() runs the whole data-parallel step inside it — counters and gradient slices all-reduced
        on a communication stream beside the backward kernels.  Otherwise the gradients are summed from here with one
        torch all_reduce over the flat fp16 gradient buffer between train_begin and train_end (layout nerf_network.h:430-443)."""
        import torch.distributed as dist
        rank, world = dist.get_rank(process_group), dist.get_world_size(process_group)
        nccl = dist.get_backend(process_group) == "nccl"
        if native is None:
            native = nccl and self.mode == TestbedMode.Nerf
        if native:
            path = None
            try:
                import nvidia
                for p in nvidia.__path__:
                    cand = os.path.join(p, "nccl", "lib", "libnccl.so.2")
                    if os.path.exists(cand):
                        path = cand.encode()
            except ImportError:
                pass
            buf = (C.c_uint8 * 128)()
            if rank == 0:
                check(lib().ngp_comm_unique_id(buf, path))
            t = torch.tensor(list(buf), dtype=torch.uint8, device="cuda" if nccl else "cpu")
            dist.broadcast(t, src=dist.get_global_rank(process_group, 0) if process_group is not None else 0, group=process_group)
            ident = (C.c_uint8 * 128)(*t.cpu().tolist())
            check(lib().ngp_testbed_comm_init(self._h, ident, rank, world, path))
            object.__setattr__(self, "_dist", None)          # train() is the plain entry point: the step is data-parallel inside
            return
        side = torch.cuda.Stream() if nccl else None
        object.__setattr__(self, "_dist", (dist, process_group, side))
        check(lib().ngp_testbed_set_shard(self._h, rank, world))

    def train(self, batch_size=None):
        n = int(batch_size or self.training_batch_size)
        if self._dist is None:
            check(lib().ngp_testbed_train(self._h, _stream(), n))
            return
        dist, pg, side = self._dist
        w = dist.get_world_size(pg)
        check(lib().ngp_testbed_train_begin(self._h, _stream(), n))
        g = self.param_buffer(3)
        if self.mode == TestbedMode.Nerf:
            # keep rays_per_batch in lock-step: the control law sees the mean sample counts.  The local counts are on the host
            # as soon as the loss kernel has run; their all-reduce goes out on a side stream, i.e. NOT behind the backward pass
            # that the training stream is still executing, and the host has the sums before the gradients are complete.
            m, mb = C.c_uint32(), C.c_uint32()
            check(lib().ngp_testbed_pending_counters(self._h, C.byref(m), C.byref(mb)))
            if side is not None:
                with torch.cuda.stream(side):
                    c = torch.tensor([m.value, mb.value], dtype=torch.int64, device=g.device)
                    dist.all_reduce(c, group=pg)
                    c = c.cpu()
            else:
                c = torch.tensor([m.value, mb.value], dtype=torch.int64)
                dist.all_reduce(c, group=pg)
            m, mb = max(int(c[0]) // w, 1), max(int(c[1]) // w, 1)
        else:
            m, mb = 1, 1
        # gradient sum: enqueued behind the backward pass; the optimizer (train_end) is ordered behind it on the stream, while
        # the next step's ray prefetch, launched by train_end on its own stream, runs beside both
        dist.all_reduce(g, group=pg)
        check(lib().ngp_testbed_train_end(self._h, _stream(), m, mb))

    def frame(self):
        if self.shall_train:
            self.train(self.training_batch_size)
        return True

    @property
    def stats(self):
        st = TrainStats()
        check(lib().ngp_testbed_stats(self._h, C.byref(st)))
        return st

    @property
    def training_step(self):
        return int(self.get_option("training_step"))      # the host's step counter: no wait for the step in flight (stats / loss do wait)

    @property
    def loss(self):
        return self.stats.loss

    # ---- buffers ----
    def param_buffer(self, which):
        ptr, n = C.c_void_p(), C.c_uint64()
        check(lib().ngp_testbed_param_buffer(self._h, which, C.byref(ptr), C.byref(n)))
        # 0 fp32 master, 1 fp16 params, 2 fp16 inference (EMA) params, 3 fp16 grads, 4 / 5 Adam first / second moments (fp32),
        # 6 per-entry step counts (u32 as int32), 7 fp32 EMA (null without an Ema optimizer)
        if not ptr.value:
            return None
        return _wrap(ptr.value, n.value, {0: torch.float32, 4: torch.float32, 5: torch.float32, 6: torch.int32, 7: torch.float32}.get(which, torch.float16))

    def density_grid(self, which=0):
        ptr, n = C.c_void_p(), C.c_uint64()
        check(lib().ngp_testbed_density_grid(self._h, which, C.byref(ptr), C.byref(n)))
        return _wrap(ptr.value, n.value // (4 if which == 0 else 1), torch.float32 if which == 0 else torch.uint8)

    def debug_buffer(self, name):
        ptr, n = C.c_void_p(), C.c_uint64()
        check(lib().ngp_testbed_debug_buffer(self._h, name.encode(), C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def debug_tensor(self, name, dtype, count):
        ptr, n = self.debug_buffer(name)
        return _wrap(ptr, count, dtype)

    # ---- inference ----
    def render(self, width=1920, height=1080, spp=1, linear=True, start_t=-1.0, end_t=-1.0, fps=30.0, shutter_fraction=1.0, *, camera_matrix=None,
               focal_length=None, screen_center=(0.5, 0.5), y0=0, y1=None, row_stride=1, out=None, return_depth=False):
        """Testbed.render(width, height, spp, linear, start_t, end_t, fps, shutter_fraction) (python_api.cu:127-205,416-425), headless: the frame
        the testbed's own camera sees (camera_matrix / fov / fov_axis / zoom / screen_center, set_camera_to_training_view ...), spp samples
        accumulated, composited over background_color with exposure / tonemap_curve, sRGB-encoded unless `linear`, ground truth overlaid when
        render_ground_truth is set; float32 [rows, W, 4].  start_t / end_t move the camera along the GUI's camera path, which this library
        does not have: anything but the default -1 raises.

        Extension (keyword only): an explicit `camera_matrix` (12 floats, mat4x3 column major, NGP convention) with `focal_length` in
        pixels renders the raw tracer output instead — linear premultiplied RGBA, no background, no tonemapping; `y0` / `y1` / `row_stride`
        restrict either form to rows y0, y0 + row_stride, ... < y1 (tile-sharded inference: a contiguous band with stride 1, or every
        world-th row starting at the rank, which balances the ranks whatever the picture shows), `out` may be a CUDA tensor."""
        if start_t >= 0.0 or end_t >= 0.0:
            raise RuntimeError("render: camera-path animation (start_t / end_t) is not part of this library (no camera path)")
        y1 = height if y1 is None else y1
        self.sync_inference_params()      # no-op unless the sharded optimizer left other ranks' EMA slices behind (then: collective)
        rows = len(range(y0, y1, int(row_stride)))
        dev_out = out is not None and isinstance(out, torch.Tensor) and out.is_cuda
        if out is None:
            out = np.empty((rows, width, 4), np.float32)
        elif tuple(out.shape) != (rows, width, 4):
            raise ValueError(f"render: out must have shape {(rows, width, 4)}")
        optr = C.c_void_p(out.data_ptr()) if isinstance(out, torch.Tensor) else out.ctypes.data_as(C.c_void_p)
        check(lib().ngp_testbed_set_option(self._h, b"render_row_stride", float(int(row_stride))))
        try:
            if camera_matrix is None:
                if return_depth:
                    raise RuntimeError("render: return_depth needs the explicit-camera form")
                check(lib().ngp_testbed_render_frame(self._h, _stream(), width, height, spp, int(bool(linear)), y0, y1, optr, int(dev_out)))
                return out
            cam = (C.c_float * 12)(*[float(x) for x in np.asarray(camera_matrix, np.float32).reshape(-1)])
            fl = (C.c_float * 2)(*focal_length)
            sc = (C.c_float * 2)(*screen_center)
            depth = np.empty((rows, width), np.float32) if (return_depth and not dev_out) else None
            dptr = depth.ctypes.data_as(C.c_void_p) if depth is not None else None
            check(lib().ngp_testbed_render(self._h, _stream(), cam, fl, sc, width, height, spp, y0, y1, optr, int(dev_out), dptr))
            return (out, depth) if return_depth else out
        finally:
            check(lib().ngp_testbed_set_option(self._h, b"render_row_stride", 1.0))

    def n_training_views(self):
        return lib().ngp_testbed_n_training_views(self._h)

    def training_view(self, i):
        """nerf.training.dataset.metadata[i] / xforms[i] as a view dict (no pixels)"""
        v = TrainingImage()
        check(lib().ngp_testbed_training_view(self._h, int(i), C.byref(v)))
        return {"res": tuple(v.res), "focal_length": tuple(v.focal_length), "principal_point": tuple(v.principal_point), "lens_mode": v.lens_mode,
                "lens_params": tuple(v.lens_params), "xform": np.array(list(v.xform), np.float32), "data_type": v.data_type, "device_pixels": v.pixels}

    def set_render_lens(self, lens_mode=0, lens_params=(0, 0, 0, 0)):
        check(lib().ngp_testbed_set_render_lens(self._h, int(lens_mode), (C.c_float * 4)(*lens_params)))

    def render_training_view(self, i, spp=1):
        """set_camera_to_training_view(i) + render at the view's own resolution, with its lens (scripts/run.py:236-240)"""
        v = self.training_view(i)
        self.set_render_lens(v["lens_mode"], v["lens_params"])
        try:
            return self.render(v["res"][0], v["res"][1], spp, camera_matrix=v["xform"], focal_length=v["focal_length"], screen_center=v["principal_point"])
        finally:
            self.set_render_lens(0)

    def training_view_pixels(self, i):
        """the device copy of training view i as float32 [H,W,4] linear premultiplied RGBA (what render_ground_truth shows)"""
        v = self.training_view(i)
        w, h = v["res"]
        if not v["device_pixels"]:
            raise RuntimeError("training view has no pixels (metadata restored from a snapshot: load the data set to get the images)")
        if v["data_type"] == 0:
            px = _wrap(v["device_pixels"], w * h * 4, torch.uint8).view(h, w, 4).float() / 255.0
            rgb = torch.where(px[..., :3] <= 0.04045, px[..., :3] / 12.92, ((px[..., :3] + 0.055) / 1.055) ** 2.4) * px[..., 3:4]
            return torch.cat([rgb, px[..., 3:4]], dim=2)
        return _wrap(v["device_pixels"], w * h * 4, torch.float32).view(h, w, 4).clone()

    def evaluate_training_views(self, view_indices=None, spp=8):
        """The evaluation loop of scripts/run.py:208-266 over the loaded views, statement for statement through the same members a user of
        the reference calls: black background, rays through pixel centres, render_min_transmittance 1e-4, the ground truth rendered with
        render_ground_truth + set_camera_to_training_view + render(w, h, 1, True), the model with render(w, h, spp, True), MSE / SSIM
        (evaluation.compute_error) on the clipped sRGB images.  Frames stay on the device.  Returns {"psnr", "min_psnr", "max_psnr",
        "psnr_avgmse", "mse", "ssim", "n"} (psnr / ssim = means of the per-view values, as run.py prints them)."""
        from .evaluation import compute_error, linear_to_srgb, mse2psnr
        idx = list(range(self.n_training_views())) if view_indices is None else list(view_indices)
        saved = (self.background_color, self.snap_to_pixel_centers, self.render_min_transmittance, self.render_ground_truth, self.camera_matrix,
                 self.relative_focal_length, self.screen_center, self.training_view_index, self.render_with_lens_distortion, self.scale)
        self.background_color = [0.0, 0.0, 0.0, 1.0]
        self.snap_to_pixel_centers = True
        self.nerf.render_min_transmittance = 1e-4
        psnrs, mses, ssims = [], [], []
        try:
            meta = self.nerf.training.dataset.metadata
            for i in idx:
                w, h = meta[i].resolution
                buf = torch.empty((2, h, w, 4), dtype=torch.float32, device="cuda")
                self.render_ground_truth = True
                self.set_camera_to_training_view(i)
                ref_image = self.render(w, h, 1, True, out=buf[0])
                self.render_ground_truth = False
                image = self.render(w, h, spp, True, out=buf[1])
                A = linear_to_srgb(image[..., :3]).clamp(0.0, 1.0)
                R = linear_to_srgb(ref_image[..., :3]).clamp(0.0, 1.0)
                mse = float(compute_error("MSE", A, R)); ssim = float(compute_error("SSIM", A, R))
                mses.append(mse); ssims.append(ssim); psnrs.append(float(mse2psnr(max(mse, 1e-12))))
        finally:
            (bg, self.snap_to_pixel_centers, self.render_min_transmittance, self.render_ground_truth, cam, rfl, sc, tv, lens, scale) = saved
            self.background_color = bg; self.camera_matrix = cam; self.relative_focal_length = rfl; self.screen_center = sc
            self.training_view_index = tv; self.render_with_lens_distortion = lens
            check(lib().ngp_testbed_set_option(self._h, b"scale", float(scale)))
        n = max(len(idx), 1)
        return {"psnr": float(np.sum(psnrs) / n), "min_psnr": float(np.min(psnrs)), "max_psnr": float(np.max(psnrs)),
                "psnr_avgmse": float(mse2psnr(max(np.sum(mses) / n, 1e-12))), "mse": float(np.sum(mses) / n), "ssim": float(np.sum(ssims) / n), "n": len(idx)}

    def compute_image_mse(self, quantize=False):
        v = C.c_float()
        check(lib().ngp_testbed_compute_image_mse(self._h, _stream(), int(quantize), C.byref(v)))
        return v.value


class _DevPtr:
    def __init__(self, ptr, count, dtype):
        item = torch.empty(0, dtype=dtype).element_size()
        typestr = {torch.float32: "<f4", torch.float16: "<f2", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (ptr, False), "version": 2, "strides": None}
        self._bytes = item * count


def _wrap(ptr, count, dtype):
    """zero-copy torch view of library-owned device memory"""
    return torch.as_tensor(_DevPtr(ptr, count, dtype), device="cuda")
