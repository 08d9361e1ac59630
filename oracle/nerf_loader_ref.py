"""ORACLE — TEST INFRASTRUCTURE ONLY.  numpy restatement of the host side of load_nerf
(/root/reference/src/nerf_loader.cu:161-258 read_lens / read_focal_length, :319-380 frame order + sharpness culling,
:405-520 scene fields, :671-697 per-frame intrinsics + pose; include/neural-graphics-primitives/nerf_loader.h:120-140
nerf_matrix_to_ngp).  May be imported only by tests/.  SI::natural::compare is a third-party header absent from the mount:
restated as "digit runs compare by value"."""
import json
import math
import os
import re

import numpy as np

F = np.float32
FORMATS = ["png", "jpg", "jpeg", "bmp", "gif", "tga", "pic", "pnm", "psd", "exr"]


def _strip_comments(text):
    return re.sub(r"/\*.*?\*/", "", re.sub(r"(?m)^\s*//.*$", "", text), flags=re.S)


def natural_key(s):
    return [(0, int(t), len(t)) if t.isdigit() else (1, t, 0) for t in re.findall(r"\d+|\D", s)]


def resolve_path(base, local):
    path = local if os.path.isabs(local) else os.path.join(base, local)
    if os.path.splitext(path)[1] == "" and not os.path.exists(path):
        for f in FORMATS:
            if os.path.exists(path + "." + f):
                return path + "." + f
    return path


def fov_to_focal_length(resolution, degrees):
    return F(0.5) * F(resolution) / F(math.tan(float(F(0.5) * F(degrees) * F(math.pi) / F(180.0))))


def read_focal_length(j, res):
    def one(r, axis):
        if axis + "_fov" in j:
            return fov_to_focal_length(r, F(j[axis + "_fov"]))
        if "fl_" + axis in j:
            return F(j["fl_" + axis])
        if "camera_angle_" + axis in j:
            return fov_to_focal_length(r, F(j["camera_angle_" + axis]) * F(180) / F(math.pi))
        return F(0)
    x, y = one(res[0], "x"), one(res[1], "y")
    if x != 0:
        return (x, y if y != 0 else x)
    if y != 0:
        return (y, y)
    return None


def read_lens(j, lens):
    mode = 0
    for name, idx in (("k1", 0), ("k2", 1), ("k3", 2), ("k4", 3), ("p1", 2), ("p2", 3)):
        if name in j:
            lens["params"][idx] = F(j[name])
            if lens["params"][idx] != 0:
                mode = 1
    if "cx" in j:
        lens["pp"][0] = F(j["cx"]) / F(j["w"])
    if "cy" in j:
        lens["pp"][1] = F(j["cy"]) / F(j["h"])
    if "rolling_shutter" in j:      # nerf_loader.cu:190-202: [A, B, C] or [A, B, C, D]
        rs = j["rolling_shutter"]
        lens["rolling_shutter"] = [F(rs[0]), F(rs[1]), F(rs[2]), F(rs[3]) if len(rs) >= 4 else F(0)]
    if mode:
        lens["mode"] = mode


def nerf_matrix_to_ngp(m34, scale, offset, from_mitsuba=False):
    """m34: rows 0-2 of transform_matrix -> NGP mat4x3 as 12 floats, column major"""
    r = np.array(m34, F)[:3, :4].copy()
    r[:, 1] *= F(-1); r[:, 2] *= F(-1)
    r[:, 3] = r[:, 3] * F(scale) + np.asarray(offset, F)
    if from_mitsuba:
        r[:, 0] *= F(-1); r[:, 2] *= F(-1)
    else:
        r = r[[1, 2, 0], :]
    return r.T.reshape(-1).copy()


def load(json_path, resolutions=None):
    """-> dict(scale, offset, aabb_scale, frames=[dict(path, xform, focal, pp, lens_mode, lens_params)])"""
    base = os.path.dirname(os.path.abspath(json_path))
    j = json.loads(_strip_comments(open(json_path).read()))
    frames = list(j["frames"])
    for f in frames:
        f["file_path"] = f["file_path"].replace("\\", "/")
    thr = F(j.get("sharpness_discard_threshold", 0.0))
    frames.sort(key=lambda f: natural_key(f["file_path"]))
    if "n_frames" in j:
        frames = frames[:int(j["n_frames"])]
    if "sharpness" in frames[0]:
        copy, frames = frames, []
        n = len(copy)
        for i in range(n):
            lo, hi = max(0, i - 3), min(i + 3, n - 1)
            mean = F(0)
            for k in range(lo, hi):
                mean = F(mean + F(copy[k].get("sharpness", 1.0)))
            with np.errstate(all="ignore"):
                mean = F(mean) / F(hi - lo)
            if os.path.exists(resolve_path(base, copy[i]["file_path"])) and copy[i].get("sharpness", 1.0) > float(F(thr * mean)):
                frames.append(copy[i])
    scale, offset = F(0.33), np.array([0.5, 0.5, 0.5], F)
    from_mitsuba = "normal_mts_args" in j
    if from_mitsuba:
        scale = F(0.66); offset = np.full(3, F(0.25) * scale, F)
    if "scale" in j:
        scale = F(j["scale"])
    lens = {"mode": 0, "params": [F(0)] * 4, "pp": [F(0.5), F(0.5)], "rolling_shutter": [F(0)] * 4}
    read_lens(j, lens)
    aabb_scale = int(j.get("aabb_scale", 1))
    if "offset" in j:
        o = j["offset"]
        offset = np.array(o if isinstance(o, list) else [o] * 3, F)
    if "aabb" in j:
        lo, hi = np.array(j["aabb"][0], F), np.array(j["aabb"][1], F)
        length = max(F(0.000001), np.abs(hi - lo).max())
        scale = F(1) / F(length)
        offset = ((hi + lo) * F(0.5)) * -scale + F(0.5)
    # depth supervision data (nerf_loader.cu:404,420-423,471-473,609-621,711): a frame's depth file is read only when the scene gives an
    # "integer_depth_scale" (> 0) and loading is enabled; the map is stored as pixel * integer_depth_scale * scene scale
    enable_depth_loading = bool(j.get("enable_depth_loading", True))
    depth_scale = F(j["integer_depth_scale"]) if "integer_depth_scale" in j else F(-1.0)
    out = []
    for i, f in enumerate(frames):
        l = {"mode": lens["mode"], "params": list(lens["params"]), "pp": list(lens["pp"]), "rolling_shutter": list(lens["rolling_shutter"])}
        read_lens(f, l)
        m = f.get("transform_matrix_start", f.get("transform_matrix"))
        m_end = f.get("transform_matrix_end", m)             # nerf_loader.cu:648-649
        rec = dict(path=resolve_path(base, f["file_path"]), xform=nerf_matrix_to_ngp(m, scale, offset, from_mitsuba), pp=l["pp"], lens_mode=l["mode"],
                   lens_params=l["params"], focal=None, xform_end=nerf_matrix_to_ngp(m_end, scale, offset, from_mitsuba), rolling_shutter=l["rolling_shutter"])
        rec["depth_path"] = ""
        if enable_depth_loading and depth_scale > 0 and "depth_path" in f:
            dp = resolve_path(base, f["depth_path"])
            if os.path.exists(dp):
                rec["depth_path"] = dp
        rec["depth_scale"] = F(depth_scale * scale)          # < 0: the view has no depth map at all
        if resolutions is not None:
            res = resolutions[i]
            fl = read_focal_length(j, res)
            fl2 = read_focal_length(f, res)
            rec["focal"] = fl2 if fl2 is not None else fl
        out.append(rec)
    return dict(scale=scale, offset=offset, aabb_scale=aabb_scale, from_mitsuba=from_mitsuba, frames=out)
