"""Testbed.render with the testbed's own view state + the output stage (python_api.cu:127-205,452-537; render_buffer.cu:232-346,533-565) and
the evaluation loop of scripts/run.py:208-266 run through the binding's public members.  The CUDA kernels of csrc/frame_output.cu are
compared with the numpy restatement oracle/frame_output_ref.py; the view-state path is compared with the explicit-camera tracer entry
point that tests/test_nerf_gpu.py::test_render_matches_oracle_and_tiles pins against the oracle tracer."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CFG = {
    "loss": {"otype": "Huber"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.5},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
N_VIEWS, RES, BATCH = 8, 64, 1 << 14
TOL = dict(rtol=2e-5, atol=2e-6)      # powf / division of the device against numpy float32: a few ulp


@pytest.fixture(scope="module")
def world(ngp):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    fo = importlib.import_module("oracle.frame_output_ref")
    views = scene.make_views(N_VIEWS, RES, RES)
    views[2]["principal_point"] = (0.45, 0.55)                     # an off-centre view: screen_center = 1 - principal_point
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(views, aabb_scale=1)
    tb.reload_network_from_json(CFG)
    for _ in range(600):      # ~22 dB / SSIM 0.86 on the training views (11 dB after 150 steps: the volume is still fog)
        tb.train(BATCH)
    return dict(pyngp=pyngp, fo=fo, views=views, tb=tb, scene=scene)


def test_default_view_state(world):
    pyngp = world["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    assert tb.fov == pytest.approx(50.625, abs=1e-4) and tb.fov_axis == 1 and tb.zoom == 1.0 and tb.scale == 1.5 and tb.exposure == 0.0
    np.testing.assert_allclose(tb.camera_matrix, [[1, 0, 0, 0.5], [0, -1, 0, 0.5], [0, 0, -1, 2.0]])      # reset_camera (testbed.cu:858-884)
    np.testing.assert_allclose(tb.look_at, [0.5, 0.5, 0.5], atol=1e-6)
    np.testing.assert_allclose(tb.background_color, [0, 0, 0, 1]); np.testing.assert_allclose(tb.screen_center, [0.5, 0.5])
    assert tb.render_mode == pyngp.RenderMode.Shade and tb.color_space == pyngp.ColorSpace.Linear and tb.tonemap_curve == pyngp.TonemapCurve.Identity
    assert not tb.render_ground_truth and not tb.render_groundtruth
    tb.fov = 40.0
    assert tb.fov == pytest.approx(40.0, abs=1e-4) and tb.fov_xy[0] == pytest.approx(40.0, abs=1e-4)
    tb.fov_xy = [30.0, 60.0]
    np.testing.assert_allclose(tb.fov_xy, [30.0, 60.0], atol=1e-4)
    tb.scale = 3.0                                                     # the eye moves back along the view axis, look_at stays
    np.testing.assert_allclose(tb.camera_matrix[:, 3], [0.5, 0.5, 3.5], atol=1e-6); np.testing.assert_allclose(tb.look_at, [0.5, 0.5, 0.5], atol=1e-6)
    tb.view_dir = [1.0, 0.0, 0.0]
    np.testing.assert_allclose(tb.view_dir, [1, 0, 0], atol=1e-6); np.testing.assert_allclose(tb.look_at, [0.5, 0.5, 0.5], atol=1e-5)
    tb.render_aabb = pyngp.BoundingBox([0.1, 0.2, 0.3], [0.9, 0.8, 0.7])
    assert np.allclose(tb.render_aabb.min, [0.1, 0.2, 0.3]) and np.allclose(tb.aabb.max, [1, 1, 1])
    tb.reset_camera()
    assert tb.fov == pytest.approx(50.625, abs=1e-4) and tb.scale == 1.5
    # set_nerf_camera_matrix: NeRF convention -> NGP (nerf_loader.h:120-140) with the default scale 0.33 / offset 0.5
    c2w = np.array([[1, 0, 0, 1.0], [0, 1, 0, 2.0], [0, 0, 1, 3.0]], np.float32)
    tb.set_nerf_camera_matrix(c2w)
    m = tb.camera_matrix
    np.testing.assert_allclose(m[:, 3], [2 * 0.33 + 0.5, 3 * 0.33 + 0.5, 1 * 0.33 + 0.5], atol=1e-6)      # cyclic axis permutation of the scaled position
    np.testing.assert_allclose(m[:, 0], [0, 0, 1], atol=1e-6); np.testing.assert_allclose(m[:, 1], [-1, 0, 0], atol=1e-6); np.testing.assert_allclose(m[:, 2], [0, -1, 0], atol=1e-6)
    with pytest.raises(RuntimeError):
        tb.render(8, 8, 1, True, start_t=0.0)


def _gt(world, i):
    v = world["views"][i]
    return v["pixels"].numpy()


@pytest.mark.parametrize("case", [
    dict(view=0, res=(RES, RES), bg=[0, 0, 0, 1], exposure=0.0, linear=True, cs=0, curve=0, alpha=1.0),
    dict(view=1, res=(40, 56), bg=[0.3, 0.5, 0.7, 1.0], exposure=0.75, linear=False, cs=0, curve=1, alpha=1.0),      # other resolution: nearest lookup, fov axis 1
    dict(view=3, res=(80, 48), bg=[0.9, 0.1, 0.4, 0.5], exposure=-1.0, linear=True, cs=1, curve=2, alpha=1.0),      # accumulated in sRGB, Hable
    dict(view=4, res=(RES, RES), bg=[1, 1, 1, 1], exposure=0.0, linear=False, cs=0, curve=3, alpha=1.0),
])
def test_ground_truth_frame_matches_restatement(world, case):
    """render_ground_truth with ground_truth_alpha 1: nothing is traced; accumulate (empty frame) -> tonemap -> overlay_image"""
    tb, fo = world["tb"], world["fo"]
    w, h = case["res"]
    tb.background_color = case["bg"]; tb.exposure = case["exposure"]; tb.color_space = case["cs"]; tb.tonemap_curve = case["curve"]
    tb.render_ground_truth = True
    tb.set_camera_to_training_view(case["view"])
    try:
        got = tb.render(w, h, 1, case["linear"])
    finally:
        tb.render_ground_truth = False; tb.background_color = [0, 0, 0, 1]; tb.exposure = 0.0; tb.color_space = 0; tb.tonemap_curve = 0
    out_cs = fo.LINEAR if case["linear"] else fo.SRGB
    surface = fo.tonemap_pass(np.zeros((h, w, 4), np.float32), case["exposure"], case["bg"], case["cs"], out_cs, case["curve"])
    want = fo.overlay_image(surface, case["alpha"], [case["exposure"]] * 3, case["bg"], out_cs, _gt(world, case["view"]), 0, case["curve"], case["cs"], 1, 1.0)
    assert got.shape == (h, w, 4) and got.dtype == np.float32
    np.testing.assert_allclose(got, want, **TOL)
    assert got[..., :3].max() > 0.05


@pytest.mark.parametrize("case", [
    dict(view=2, spp=1, bg=[0, 0, 0, 1], exposure=0.0, linear=True, cs=0, curve=0),
    dict(view=2, spp=3, bg=[0.2, 0.6, 0.4, 1.0], exposure=0.5, linear=False, cs=0, curve=1),
    dict(view=5, spp=2, bg=[0.2, 0.6, 0.4, 0.25], exposure=0.0, linear=True, cs=0, curve=3),
])
def test_view_state_render_equals_explicit_camera_plus_output_stage(world, ref, case):
    """set_camera_to_training_view(i) + render(w, h, spp, linear) == the tracer with that view's camera / focal length / principal point
    (explicit-camera entry point, pinned against the oracle tracer elsewhere) followed by the restated accumulate + tonemap"""
    tb, fo = world["tb"], world["fo"]
    v = world["views"][case["view"]]
    w, h = 56, 72                                                      # not the view's resolution: focal length follows the fov axis (y)
    tb.background_color = case["bg"]; tb.exposure = case["exposure"]; tb.color_space = case["cs"]; tb.tonemap_curve = case["curve"]
    tb.set_camera_to_training_view(case["view"])
    try:
        got = tb.render(w, h, case["spp"], case["linear"])
        assert tb.nerf.training.view == case["view"] and tb.render_with_lens_distortion
        # set_camera_to_training_view takes the pose through get_xform_given_rolling_shutter(.., uv = 0.5, t = 0) (testbed.cu:843): the quaternion round trip
        view_cam = ref.xform_given_rolling_shutter(v["xform"])
        assert np.array_equal(tb.camera_matrix, view_cam.reshape(4, 3).T)
        np.testing.assert_allclose(tb.camera_matrix, np.asarray(v["xform"], np.float32).reshape(4, 3).T, atol=2e-6)
    finally:
        tb.background_color = [0, 0, 0, 1]; tb.exposure = 0.0; tb.color_space = 0; tb.tonemap_curve = 0
    meta = {"focal_length": v["focal_length"], "res": (RES, RES), "principal_point": v["principal_point"], "xform": v["xform"]}
    cam, focal, centre = fo.view_of_training_image(meta, w, h)
    cam = view_cam
    raw = tb.render(w, h, case["spp"], camera_matrix=cam, focal_length=[float(f) for f in focal], screen_center=[float(c) for c in centre])
    want = fo.tonemap_pass(raw, case["exposure"], case["bg"], case["cs"], fo.LINEAR if case["linear"] else fo.SRGB, case["curve"])
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-5)
    assert raw[..., 3].max() > 0.9


def test_fov_and_camera_matrix_drive_the_frame(world):
    tb = world["tb"]
    v = world["views"][6]
    w, h = 48, 32
    tb.reset_camera()
    tb.render_with_lens_distortion = False
    tb.camera_matrix = np.asarray(v["xform"], np.float32).reshape(4, 3).T
    tb.fov = 35.0
    got = tb.render(w, h, 1, True)
    f = float(np.float32(tb.relative_focal_length[1]) * np.float32(h))               # calc_focal_length: relative focal length x resolution[fov_axis]
    assert f == pytest.approx(0.5 * h / np.tan(0.5 * np.radians(35.0)), rel=1e-6)
    raw = tb.render(w, h, 1, camera_matrix=v["xform"], focal_length=(f, f))
    np.testing.assert_allclose(got[..., :3], raw[..., :3], rtol=1e-5, atol=1e-6)      # black opaque background: colours unchanged
    np.testing.assert_allclose(got[..., 3], 1.0, rtol=1e-6)
    tb.fov_axis = 0
    got_x = tb.render(w, h, 1, True)
    fx = float(np.float32(tb.relative_focal_length[0]) * np.float32(w))
    raw_x = tb.render(w, h, 1, camera_matrix=v["xform"], focal_length=(fx, fx))
    np.testing.assert_allclose(got_x[..., :3], raw_x[..., :3], rtol=1e-5, atol=1e-6)
    tb.fov_axis = 1


def test_render_modes(world):
    tb, pyngp = world["tb"], world["pyngp"]
    tb.set_camera_to_training_view(0)
    tb.background_color = [0, 0, 0, 0]
    try:
        shade = tb.render(RES, RES, 1, True)
        tb.render_mode = pyngp.RenderMode.Positions
        pos = tb.render(RES, RES, 1, True)
        tb.render_mode = pyngp.RenderMode.Depth
        depth = tb.render(RES, RES, 1, True)
        tb.render_mode = pyngp.RenderMode.Cost
        cost = tb.render(RES, RES, 1, True)
        tb.render_mode = pyngp.RenderMode.AO
        ao = tb.render(RES, RES, 1, True)
        tb.render_mode = pyngp.RenderMode.Normals
        with pytest.raises(Exception, match="Normals"):
            tb.render(RES, RES, 1, True)
    finally:
        tb.render_mode = pyngp.RenderMode.Shade; tb.background_color = [0, 0, 0, 1]
    hit = shade[..., 3] > 0.9
    assert hit.mean() > 0.05
    np.testing.assert_allclose(pos[..., 3], shade[..., 3], rtol=1e-6)                  # same alpha whatever is composited
    p = pos[hit][:, :3] / pos[hit][:, 3:4]
    assert p.min() > 0.2 and p.max() < 0.8                                             # (pos - 0.5) / 2 + 0.5 of points inside the unit cube
    d = depth[hit][:, 0] / depth[hit][:, 3]
    dist_cam = np.linalg.norm(np.asarray(world["views"][0]["xform"], np.float32)[9:12] - 0.5)
    assert np.all(np.abs(d * 0.33 - dist_cam) < 0.9)                                    # z depth in dataset units (depth_scale = 1 / 0.33), surface near the centre
    assert np.allclose(depth[..., 0], depth[..., 1]) and np.allclose(ao[..., 0], ao[..., 2])
    assert np.all(cost[hit][:, 3] == 1.0) and cost[hit][:, 0].min() > 0 and np.allclose(cost[hit][:, 0] * 128, np.round(cost[hit][:, 0] * 128))


def test_run_py_evaluation_loop_through_the_binding(world):
    """scripts/run.py:208-266, statement for statement, on the binding's members; numbers against evaluate_training_views (same loop kept on the device)"""
    tb, pyngp = world["tb"], world["pyngp"]
    ev_mod = importlib.import_module("instant-ngp_b200.evaluation")
    compute_error, linear_to_srgb, mse2psnr = ev_mod.compute_error, ev_mod.linear_to_srgb, ev_mod.mse2psnr
    testbed = tb
    totmse = totpsnr = totssim = totcount = 0
    minpsnr, maxpsnr = 1000, 0
    testbed.background_color = [0.0, 0.0, 0.0, 1.0]
    testbed.snap_to_pixel_centers = True
    spp = 8
    testbed.nerf.render_min_transmittance = 1e-4
    testbed.shall_train = False
    for i in range(testbed.nerf.training.dataset.n_images):
        resolution = testbed.nerf.training.dataset.metadata[i].resolution
        testbed.render_ground_truth = True
        testbed.set_camera_to_training_view(i)
        ref_image = testbed.render(resolution[0], resolution[1], 1, True)
        testbed.render_ground_truth = False
        image = testbed.render(resolution[0], resolution[1], spp, True)
        A = np.clip(linear_to_srgb(image[..., :3]), 0.0, 1.0)
        R = np.clip(linear_to_srgb(ref_image[..., :3]), 0.0, 1.0)
        mse = float(compute_error("MSE", A, R))
        ssim = float(compute_error("SSIM", A, R))
        totssim += ssim; totmse += mse
        psnr = mse2psnr(mse)
        totpsnr += psnr
        minpsnr = psnr if psnr < minpsnr else minpsnr
        maxpsnr = psnr if psnr > maxpsnr else maxpsnr
        totcount += 1
    psnr, ssim = totpsnr / totcount, totssim / totcount
    testbed.snap_to_pixel_centers = False; testbed.nerf.render_min_transmittance = 0.01; testbed.shall_train = True
    assert totcount == N_VIEWS and 18 < psnr < 60 and 0.5 < ssim <= 1.0, (psnr, ssim)
    ev = tb.evaluate_training_views(None, spp=8)
    assert ev["psnr"] == pytest.approx(psnr, abs=2e-3) and ev["ssim"] == pytest.approx(ssim, abs=1e-4), (ev, psnr, ssim)
    assert ev["min_psnr"] == pytest.approx(minpsnr, abs=2e-3) and ev["max_psnr"] == pytest.approx(maxpsnr, abs=2e-3)
    assert ev["psnr_avgmse"] == pytest.approx(mse2psnr(totmse / totcount), abs=2e-3)
    assert tb.render_min_transmittance == pytest.approx(0.01) and not tb.snap_to_pixel_centers and not tb.render_ground_truth      # knobs restored
