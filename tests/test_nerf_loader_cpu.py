"""Host side of the NeRF dataset loader (csrc/nerf_dataset.cu behind ngp_nerf_dataset_*) against the numpy restatement of
load_nerf in oracle/nerf_loader_ref.py: frame order, culling, scale / offset, intrinsics, lens, NeRF -> NGP poses.
No device needed.  The fox scene of the reference mount is checked when the mount is present (it is not on the GPU box)."""
import ctypes as C
import importlib
import json
import os

import numpy as np
import pytest

from oracle import nerf_loader_ref as loader_ref

FOX = "/root/reference/data/nerf/fox/transforms.json"


@pytest.fixture(scope="module")
def L(ngp):
    return importlib.import_module("instant-ngp_b200._lib")


def _load_cpp(L, path, resolutions):
    lib = L.lib()
    d = C.c_void_p()
    L.check(lib.ngp_nerf_dataset_open(os.fspath(path).encode(), C.byref(d)))
    try:
        n = lib.ngp_nerf_dataset_n_images(d)
        buf = C.create_string_buffer(4096)
        frames = []
        for i in range(n):
            L.check(lib.ngp_nerf_dataset_image_path(d, i, buf, 4096))
            w, h = resolutions[i] if isinstance(resolutions, list) else resolutions
            px = np.zeros((h, w, 4), np.uint8)
            L.check(lib.ngp_nerf_dataset_set_image(d, i, px.ctypes.data_as(C.c_void_p), 0, w, h))
            v = L.TrainingImage()
            L.check(lib.ngp_nerf_dataset_view(d, i, C.byref(v)))
            xe, rs = np.zeros(12, np.float32), np.zeros(4, np.float32)
            L.check(lib.ngp_nerf_dataset_view_motion(d, i, xe.ctypes.data_as(C.c_void_p), rs.ctypes.data_as(C.c_void_p)))
            frames.append(dict(path=buf.value.decode(), xform=np.array(list(v.xform), np.float32), focal=tuple(v.focal_length), pp=tuple(v.principal_point),
                               lens_mode=v.lens_mode, lens_params=tuple(v.lens_params), res=tuple(v.res), xform_end=xe, rolling_shutter=rs))
            L.check(lib.ngp_nerf_dataset_depth_path(d, i, buf, 4096))
            frames[-1]["depth_path"] = buf.value.decode()
        a, s, o, m = C.c_int32(), C.c_float(), (C.c_float * 3)(), C.c_int32()
        L.check(lib.ngp_nerf_dataset_info(d, C.byref(a), C.byref(s), o, C.byref(m)))
        return dict(aabb_scale=a.value, scale=np.float32(s.value), offset=np.array(list(o), np.float32), frames=frames)
    finally:
        lib.ngp_nerf_dataset_close(d)


def _compare(cpp, ref, exact_focal=True):
    assert cpp["aabb_scale"] == ref["aabb_scale"] and cpp["scale"] == ref["scale"] and np.array_equal(cpp["offset"], ref["offset"])
    assert [f["path"] for f in cpp["frames"]] == [f["path"] for f in ref["frames"]]
    for a, b in zip(cpp["frames"], ref["frames"]):
        assert np.array_equal(a["xform"], b["xform"]), (a["path"], a["xform"], b["xform"])         # bit-exact poses
        assert np.array_equal(a["xform_end"], b["xform_end"]) and np.array_equal(a["rolling_shutter"], np.float32(b["rolling_shutter"]))
        assert a["lens_mode"] == b["lens_mode"] and np.array_equal(np.float32(a["lens_params"]), np.float32(b["lens_params"]))
        assert np.array_equal(np.float32(a["pp"]), np.float32(b["pp"]))
        assert a["depth_path"] == b["depth_path"]
        if exact_focal:
            assert np.array_equal(np.float32(a["focal"]), np.float32(b["focal"]))
        else:
            np.testing.assert_allclose(np.float32(a["focal"]), np.float32(b["focal"]), rtol=1e-6)


def _write_scene(tmp_path, top, frames, missing=()):
    import cv2
    for f in frames:
        p = f["file_path"]
        if p in missing:
            continue
        full = tmp_path / (p if os.path.splitext(p)[1] else p + ".png")
        full.parent.mkdir(parents=True, exist_ok=True)
        cv2.imwrite(str(full), np.zeros((6, 8, 3), np.uint8))
    path = tmp_path / "transforms.json"
    path.write_text("// written by the test\n" + json.dumps({**top, "frames": frames}))
    return path


def _pose(rng):
    m = np.eye(4)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    m[:3, :3] = q; m[:3, 3] = rng.normal(size=3) * 3
    return m.tolist()


def test_blender_style_scene(L, tmp_path):
    rng = np.random.default_rng(1)
    names = ["./train/r_10", "./train/r_9", "./train/r_1", "./train/r_100", "./train/r_02"]
    path = _write_scene(tmp_path, {"camera_angle_x": 0.6911112070083618}, [{"file_path": n, "rotation": 0.01, "transform_matrix": _pose(rng)} for n in names])
    cpp, ref = _load_cpp(L, path, (8, 6)), loader_ref.load(path, [(8, 6)] * 5)
    assert [os.path.basename(f["path"]) for f in cpp["frames"]] == ["r_1.png", "r_02.png", "r_9.png", "r_10.png", "r_100.png"]    # natural order
    _compare(cpp, ref, exact_focal=False)
    assert cpp["aabb_scale"] == 1 and cpp["frames"][0]["lens_mode"] == 0 and cpp["frames"][0]["res"] == (8, 6)
    assert abs(cpp["frames"][0]["focal"][0] - 0.5 * 8 / np.tan(0.5 * 0.6911112070083618)) < 1e-4


def test_colmap_style_scene_with_culling_and_overrides(L, tmp_path):
    rng = np.random.default_rng(2)
    top = {"fl_x": 1375.52, "fl_y": 1374.49, "k1": 0.0578421, "k2": -0.0805099, "p1": -0.000980296, "p2": 0.00015575, "cx": 554.558, "cy": 965.268,
           "w": 1080.0, "h": 1920.0, "aabb_scale": 8, "sharpness_discard_threshold": 0.8, "scale": 0.25, "offset": [0.4, 0.5, 0.6]}
    frames = [{"file_path": f"images\\{i:04d}.png", "sharpness": float(s), "transform_matrix": _pose(rng)}
              for i, s in enumerate([30, 31, 5, 29, 33, 28, 2, 35, 30, 31], start=1)]
    frames[3]["fl_x"] = 1400.0                                            # per-frame override of one focal axis
    frames[4].update({"k1": 0.1, "cx": 500.0, "w": 1000.0})               # per-frame lens override
    rng.shuffle(frames)
    path = _write_scene(tmp_path, top, [{**f, "file_path": f["file_path"].replace("\\", "/")} for f in frames], missing={"images/0009.png"})
    path.write_text(json.dumps({**top, "frames": frames}))                # keep the Windows separators in the file
    res = [(1080, 1920)] * 10
    cpp, ref = _load_cpp(L, path, (1080, 1920)), loader_ref.load(path, res)
    _compare(cpp, ref)
    kept = [os.path.basename(f["path"]) for f in cpp["frames"]]
    assert "0003.png" not in kept and "0007.png" not in kept and "0009.png" not in kept and len(kept) == 7     # two blurry, one missing
    assert cpp["aabb_scale"] == 8 and cpp["scale"] == np.float32(0.25)
    by = {os.path.basename(f["path"]): f for f in cpp["frames"]}
    assert by["0004.png"]["focal"] == (1400.0, 1400.0)        # a lone fl_x sets both axes (read_focal_length: focal_length = vec2(x_fl))
    assert by["0001.png"]["focal"] == (np.float32(1375.52), np.float32(1374.49))
    assert by["0005.png"]["lens_params"][0] == np.float32(0.1) and by["0005.png"]["pp"][0] == np.float32(500.0) / np.float32(1000.0)
    assert by["0001.png"]["lens_mode"] == 1 and by["0001.png"]["pp"] == (np.float32(554.558) / np.float32(1080.0), np.float32(965.268) / np.float32(1920.0))


def test_aabb_fit_and_errors(L, tmp_path):
    rng = np.random.default_rng(3)
    path = _write_scene(tmp_path, {"x_fov": 40.0, "aabb": [[-1, -2, -3], [3, 2, 1]], "n_frames": 2},
                        [{"file_path": f"f{i}.png", "transform_matrix": _pose(rng)} for i in range(4)])
    cpp, ref = _load_cpp(L, path, (8, 6)), loader_ref.load(path, [(8, 6)] * 2)
    _compare(cpp, ref, exact_focal=False)
    assert len(cpp["frames"]) == 2 and cpp["scale"] == np.float32(0.25)
    lib = L.lib(); d = C.c_void_p()
    assert lib.ngp_nerf_dataset_open(str(tmp_path / "nope.json").encode(), C.byref(d)) != 0 and b"does not exist" in lib.ngp_last_error()
    (tmp_path / "empty.json").write_text('{"frames": []}')
    assert lib.ngp_nerf_dataset_open(str(tmp_path / "empty.json").encode(), C.byref(d)) != 0 and b"No training images" in lib.ngp_last_error()
    (tmp_path / "nofov.json").write_text(json.dumps({"frames": [{"file_path": "f0.png", "transform_matrix": _pose(rng)}]}))
    L.check(lib.ngp_nerf_dataset_open(str(tmp_path / "nofov.json").encode(), C.byref(d)))
    px = np.zeros((6, 8, 4), np.uint8)
    assert lib.ngp_nerf_dataset_set_image(d, 0, px.ctypes.data_as(C.c_void_p), 0, 8, 6) != 0 and b"Couldn't read fov" in lib.ngp_last_error()
    lib.ngp_nerf_dataset_close(d)
    # per-view latent codes would widen the network input: refused, not ignored
    (tmp_path / "latents.json").write_text(json.dumps({"x_fov": 40.0, "n_extra_learnable_dims": 4, "frames": [{"file_path": "f0.png", "transform_matrix": _pose(rng)}]}))
    assert lib.ngp_nerf_dataset_open(str(tmp_path / "latents.json").encode(), C.byref(d)) != 0 and b"n_extra_learnable_dims" in lib.ngp_last_error()
    (tmp_path / "nolatents.json").write_text(json.dumps({"x_fov": 40.0, "n_extra_learnable_dims": 0, "frames": [{"file_path": "f0.png", "transform_matrix": _pose(rng)}]}))
    L.check(lib.ngp_nerf_dataset_open(str(tmp_path / "nolatents.json").encode(), C.byref(d)))
    lib.ngp_nerf_dataset_close(d)


@pytest.mark.skipif(not os.path.exists(FOX), reason="reference mount not present (GPU box)")
def test_fox_scene_of_the_reference_mount(L):
    n_all = len(json.load(open(FOX))["frames"])
    ref = loader_ref.load(FOX)
    res = [(1080, 1920)] * len(ref["frames"])
    ref = loader_ref.load(FOX, res)
    cpp = _load_cpp(L, FOX, (1080, 1920))
    _compare(cpp, ref)
    assert n_all == 67 and len(cpp["frames"]) == 50 and cpp["aabb_scale"] == 8      # frames whose JPEG is absent are dropped
    assert all(os.path.exists(f["path"]) for f in cpp["frames"])


def test_transparency_rules_and_multiple_json_files(L, tmp_path):
    """convert_rgba32 (nerf_loader.cu:49-71): white / black pixels become transparent when the scene asks for it; several json files
    describe one scene (load_nerf takes a vector of paths), frames are appended file by file."""
    rng = np.random.default_rng(4)
    a = tmp_path / "a"; b = tmp_path / "b"; a.mkdir(); b.mkdir()
    pa = _write_scene(a, {"fl_x": 50.0, "white_transparent": True, "aabb_scale": 2}, [{"file_path": f"w{i}.png", "transform_matrix": _pose(rng)} for i in range(2)])
    pb = _write_scene(b, {"fl_x": 60.0, "black_transparent": True}, [{"file_path": f"k{i}.png", "transform_matrix": _pose(rng)} for i in range(3)])
    lib = L.lib(); d = C.c_void_p()
    L.check(lib.ngp_nerf_dataset_open(f"{pa}\n{pb}".encode(), C.byref(d)))
    try:
        assert lib.ngp_nerf_dataset_n_images(d) == 5
        px = np.zeros((2, 3, 4), np.uint8); px[..., 3] = 255
        px[0, 0, :3] = 255; px[0, 1, :3] = (255, 255, 254); px[1, 0, :3] = 0; px[1, 1, :3] = (0, 1, 0)
        buf = C.create_string_buffer(4096)
        names = []
        for i in range(5):
            L.check(lib.ngp_nerf_dataset_image_path(d, i, buf, 4096)); names.append(os.path.basename(buf.value.decode()))
            L.check(lib.ngp_nerf_dataset_set_image(d, i, px.ctypes.data_as(C.c_void_p), 0, 3, 2))
        assert names == ["w0.png", "w1.png", "k0.png", "k1.png", "k2.png"]
        v = L.TrainingImage()
        L.check(lib.ngp_nerf_dataset_view(d, 4, C.byref(v)))
        got = np.ctypeslib.as_array(C.cast(v.pixels, C.POINTER(C.c_uint8)), shape=(2, 3, 4)).copy()
        # a flag stays set for the files that follow (the reference keeps `info` across the files): the second file's frames obey both
        assert got[0, 0, 3] == 0 and got[1, 0, 3] == 0                    # pure white, pure black -> transparent
        assert got[0, 1, 3] == 255 and got[1, 1, 3] == 255                # anything else keeps its alpha
        assert np.array_equal(got[..., :3], px[..., :3])
        assert tuple(v.focal_length) == (60.0, 60.0)                      # the second file's intrinsics for its own frames
        L.check(lib.ngp_nerf_dataset_view(d, 0, C.byref(v)))
        assert tuple(v.focal_length) == (50.0, 50.0)
        first = np.ctypeslib.as_array(C.cast(v.pixels, C.POINTER(C.c_uint8)), shape=(2, 3, 4))
        assert first[0, 0, 3] == 0 and first[1, 0, 3] == 255              # the first file only knows white_transparent
        a_s = C.c_int32(); L.check(lib.ngp_nerf_dataset_info(d, C.byref(a_s), None, None, None)); assert a_s.value == 2
    finally:
        lib.ngp_nerf_dataset_close(d)


def test_rolling_shutter_and_start_end_poses(L, tmp_path):
    """"rolling_shutter" ([A, B, C] or [A, B, C, D], scene level and per frame) and "transform_matrix_start" / "_end" (nerf_loader.cu:190-202,648-649,676-680)"""
    rng = np.random.default_rng(5)
    frames = [{"file_path": f"f{i}", "transform_matrix": _pose(rng)} for i in range(4)]
    frames[1] = {"file_path": "f1", "transform_matrix_start": _pose(rng), "transform_matrix_end": _pose(rng)}
    frames[2]["rolling_shutter"] = [0.1, 0.0, 0.5, 0.25]                 # per-frame override with a motion-blur term
    frames[3] = {"file_path": "f3", "transform_matrix_start": _pose(rng)}   # a start pose alone: the end pose is the start pose
    path = _write_scene(tmp_path, {"camera_angle_x": 0.7, "rolling_shutter": [0.0, 0.0, 1.0]}, frames)
    cpp, ref = _load_cpp(L, path, (8, 6)), loader_ref.load(path, [(8, 6)] * 4)
    _compare(cpp, ref, exact_focal=False)
    f = cpp["frames"]
    assert np.array_equal(f[0]["rolling_shutter"], np.float32([0, 0, 1, 0])) and np.array_equal(f[0]["xform_end"], f[0]["xform"])
    assert not np.array_equal(f[1]["xform_end"], f[1]["xform"]) and np.array_equal(f[1]["rolling_shutter"], np.float32([0, 0, 1, 0]))
    assert np.array_equal(f[2]["rolling_shutter"], np.float32([0.1, 0.0, 0.5, 0.25]))
    assert np.array_equal(f[3]["xform_end"], f[3]["xform"])


def test_depth_paths(L, tmp_path):
    """"integer_depth_scale", "enable_depth_loading" and per-frame "depth_path" (nerf_loader.cu:404,420-423,471-473,609-621): a depth file is named
    only for frames that have one on disk, and only when the scene carries a positive integer depth scale"""
    import cv2
    rng = np.random.default_rng(9)
    frames = [{"file_path": f"f{i}", "transform_matrix": _pose(rng)} for i in range(4)]
    (tmp_path / "depth").mkdir()
    for i in (0, 1, 3):
        frames[i]["depth_path"] = f"depth/d{i}.png" if i != 1 else "depth/d1"          # without extension: searched like image paths
        if i != 3:                                                                        # frame 3 names a file that is not there
            cv2.imwrite(str(tmp_path / "depth" / f"d{i}.png"), np.zeros((6, 8), np.uint16))
    top = {"camera_angle_x": 0.7, "integer_depth_scale": 0.001}
    path = _write_scene(tmp_path, top, frames)
    cpp, ref = _load_cpp(L, path, (8, 6)), loader_ref.load(path, [(8, 6)] * 4)
    _compare(cpp, ref, exact_focal=False)
    got = [os.path.basename(f["depth_path"]) for f in cpp["frames"]]
    assert got == ["d0.png", "d1.png", "", ""]
    assert ref["frames"][0]["depth_scale"] == np.float32(np.float32(0.001) * np.float32(0.33))
    for top2 in ({"camera_angle_x": 0.7}, {"camera_angle_x": 0.7, "integer_depth_scale": 0.001, "enable_depth_loading": False},
                 {"camera_angle_x": 0.7, "integer_depth_scale": 0.0}):
        path = _write_scene(tmp_path, top2, frames)
        cpp, ref = _load_cpp(L, path, (8, 6)), loader_ref.load(path, [(8, 6)] * 4)
        _compare(cpp, ref, exact_focal=False)
        assert all(f["depth_path"] == "" for f in cpp["frames"])
    # a depth map must have the resolution of its image
    lib = L.lib()
    d = C.c_void_p()
    path = _write_scene(tmp_path, top, frames)
    L.check(lib.ngp_nerf_dataset_open(os.fspath(path).encode(), C.byref(d)))
    try:
        px = np.zeros((6, 8, 4), np.uint8)
        L.check(lib.ngp_nerf_dataset_set_image(d, 0, px.ctypes.data_as(C.c_void_p), 0, 8, 6))
        bad = np.zeros((5, 8), np.uint16)
        assert lib.ngp_nerf_dataset_set_depth(d, 0, bad.ctypes.data_as(C.c_void_p), 8, 5) != 0
        assert "wrong resolution" in lib.ngp_last_error().decode()
        good = np.zeros((6, 8), np.uint16)
        L.check(lib.ngp_nerf_dataset_set_depth(d, 0, good.ctypes.data_as(C.c_void_p), 8, 6))
    finally:
        lib.ngp_nerf_dataset_close(d)


def test_host_mask_scan(L):
    """The test set_training_image applies to an RGBA8 view on the host (mask colour 0x00FF00FF of read_rgba, common_device.cuh:885-913): the
    eight-stream scan finds a single masked pixel wherever it sits (stream borders and the scalar tail included), at small and odd sizes too."""
    lib = L.lib()

    def has(px):
        out = C.c_int32(-1)
        L.check(lib.ngp_host_has_mask_pixels(px.ctypes.data_as(C.c_void_p), px.size, C.byref(out)))
        return out.value

    rng = np.random.default_rng(11)
    for n in (0, 1, 7, 1000, (1 << 17) - 1, 1 << 17, 800 * 800, 800 * 800 + 3):
        px = rng.integers(0, 1 << 32, size=max(n, 1), dtype=np.uint32)[:n]
        px[px == 0x00FF00FF] = 0
        assert has(px) == 0
        if n == 0:
            continue
        parts = [0, n - 1, n // 2] + [min(n - 1, max(0, (n + k - 1) // k * j + d)) for k in range(2, 9) for j in range(1, k) for d in (-1, 0)]
        for i in parts:
            q = px.copy(); q[i] = 0x00FF00FF
            assert has(q) == 1, (n, i)
        near = px.copy(); near[n // 3] = 0x01FF00FF                  # alpha 1: not the mask colour
        assert has(near) == 0
