"""Error-map importance sampling (nerf.training.sample_focal_plane_proportional_to_error /
sample_image_proportional_to_error, python_api.cu:624-625): the loss kernel splats every ray's mean loss into a per-view error map
(testbed_nerf.cu:1869-1900), every n_steps_between_error_map_updates steps the map becomes per-row column CDFs, per-view row CDFs and a view CDF
(:2354-2409,3700-3751), and the sampler then draws half of the pixels / all of the views from them (:1232-1338) while the loss divides the reported
loss by the pdfs (:1846).  Each stage against the oracle, in the step right after a CDF update (the first step of a fresh error-map window)."""
import importlib

import numpy as np
import pytest
import torch

from .test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(ngp, ref):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    images, keep = ref.images_from_views(views)
    return dict(pyngp=pyngp, scene=scene, views=views, images=images, keep=keep)


def _buffers(tb):
    ex, ey = int(tb.get_option("error_map_width")), int(tb.get_option("error_map_height"))
    cx, cy = int(tb.get_option("error_cdf_width")), int(tb.get_option("error_cdf_height"))
    emap = tb.debug_tensor("error_map", torch.float32, N_VIEWS * ex * ey).cpu().numpy().reshape(N_VIEWS, ey, ex).copy()
    out = dict(error_map=emap, map_res=(ex, ey), cdf_res=(cx, cy))
    if cx:
        out["cdf_x_cond_y"] = tb.debug_tensor("cdf_x_cond_y", torch.float32, N_VIEWS * cx * cy).cpu().numpy().reshape(N_VIEWS, cy, cx).copy()
        out["cdf_y"] = tb.debug_tensor("cdf_y", torch.float32, N_VIEWS * cy).cpu().numpy().reshape(N_VIEWS, cy).copy()
        out["cdf_img"] = tb.debug_tensor("cdf_img", torch.float32, N_VIEWS).cpu().numpy().copy()
    return out


@pytest.mark.parametrize("lanes,focal_plane,image", [(1, True, True), (8, True, False), (0, False, True), (32, True, True)])
def test_error_map_cdfs_and_sampling(env, ref, lanes, focal_plane, image):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tr = tb.nerf.training
    assert tr.sample_focal_plane_proportional_to_error is False and tr.sample_image_proportional_to_error is False
    assert tr.n_steps_between_error_map_updates == 128
    tr.sample_focal_plane_proportional_to_error = focal_plane
    tr.sample_image_proportional_to_error = image
    tr.n_steps_between_error_map_updates = 4
    for _ in range(3):
        tb.train(BATCH)
    assert int(tb.get_option("error_cdf_width")) == 0                       # no CDFs yet: the first window samples uniformly
    tb.train(BATCH)                                                          # 4th step: the window closes, CDFs are built
    b = _buffers(tb)
    assert tr.n_steps_between_error_map_updates == 6                         # the window grows by 1.5x
    cx, cy = b["cdf_res"]
    assert cx == b["map_res"][0] and cy == b["map_res"][1] and 4 <= cx <= RES
    # --- CDFs from the accumulated map ---
    assert b["error_map"].sum() > 0 and (b["error_map"] >= 0).all()
    wxy, wy, wimg = ref.construct_error_cdfs(b["error_map"], N_VIEWS, cy, cx)
    np.testing.assert_allclose(b["cdf_x_cond_y"], wxy, rtol=0, atol=2e-6)
    np.testing.assert_allclose(b["cdf_y"], wy, rtol=0, atol=2e-6)
    np.testing.assert_allclose(b["cdf_img"], wimg, rtol=0, atol=2e-6)
    assert np.all(np.diff(b["cdf_img"]) > 0) and abs(b["cdf_img"][-1] - 1.0) < 1e-5
    assert np.all(np.diff(b["cdf_x_cond_y"], axis=2) > 0) and np.allclose(b["cdf_x_cond_y"][:, :, -1], 1.0, atol=1e-5)

    # --- the next step draws from them: rays, samples, loss, the new window's error map ---
    st = tb.stats
    rpb, mb, step = st.rays_per_batch, st.measured_batch_size_before_compaction, st.training_step
    max_inf = BATCH * 16 if mb == 0 else (min(mb, BATCH * 16) + 255) // 256 * 256
    tb.train(BATCH)
    after = _buffers(tb)
    ex, ey = after["map_res"]
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    bf = tb.density_grid(1).cpu().numpy()
    skip = 1 + step * (1 << 32)
    emap = np.zeros((N_VIEWS, ey, ex), np.float32)
    ref.set_error_map(b["cdf_x_cond_y"] if focal_plane else None, b["cdf_y"] if focal_plane else None, b["cdf_img"] if image else None, (cx, cy), emap, (ex, ey))
    try:
        o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
        c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
        assert c[0] == o["ray_counter"] and c[1] == o["numsteps_counter"]
        nr = int(c[0]); n_written = int(o["numsteps"][:nr, 0].sum())
        assert np.array_equal(tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6), o["rays"][:nr])
        gcoords = tb.debug_tensor("coords", torch.float32, n_written * 7).cpu().numpy().reshape(n_written, 7)
        assert np.array_equal(gcoords, o["coords"][:n_written])
        mlp_out = tb.debug_tensor("mlp_out", torch.float16, n_written * 4).cpu().numpy().reshape(n_written, 4)
        mean_density = float(tb.debug_tensor("mean_density", torch.float32, 1).cpu().numpy()[0])
        cfg = ref.LossCfg(4, 2, 3, 1, 0, 1, (ref.f32 * 3)(0, 0, 0), 128.0, 0.1, mean_density)
        lo = ref.compute_loss(rpb, aabb, 1337, skip, BATCH, nr, cfg, env["images"], mlp_out, o["ray_indices"], o["rays"], o["numsteps"], o["coords"])
    finally:
        ref.set_error_map()
    # the same step sampled uniformly is another set of rays
    plain = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
    assert not np.array_equal(plain["rays"][:min(nr, int(plain["ray_counter"]))], o["rays"][:min(nr, int(plain["ray_counter"]))])
    assert c[2] == lo["compacted_counter"]
    nc = int(min(c[2], BATCH))
    gdl = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    assert np.abs(gdl[:nc] - rdl).max() <= 2e-3 * np.abs(rdl).max() + 1e-7
    gl = tb.debug_tensor("loss", torch.float32, rpb).cpu().numpy()
    np.testing.assert_allclose(gl[:nr], lo["loss"][:nr], rtol=1e-3, atol=1e-9)           # divided by the view and pixel pdfs
    # the error map of the new window holds exactly this step's rays
    assert emap.sum() > 0
    np.testing.assert_allclose(after["error_map"], emap, rtol=1e-3, atol=1e-6 * max(1.0, float(emap.max())))


def test_training_with_error_map_sampling_converges_and_can_be_switched_off(env):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.nerf.training.sample_focal_plane_proportional_to_error = True
    tb.nerf.training.sample_image_proportional_to_error = True
    tb.nerf.training.n_steps_between_error_map_updates = 16
    twin = pyngp.Testbed(pyngp.TestbedMode.Nerf)                             # the same training with uniform sampling
    twin.set_nerf_dataset(env["views"], aabb_scale=1)
    twin.reload_network_from_json(SMALL_CFG)
    for _ in range(17):
        tb.train(BATCH)
    first = tb.loss
    for _ in range(800):
        tb.train(BATCH); twin.train(BATCH)
    # (the reported loss is one 16th-step reading, weighted by 1 / pdf: noisy; it must have come down and be in the range of the uniform run's)
    assert tb.loss < first and tb.loss < 3.0 * twin.loss, (first, tb.loss, twin.loss)
    v = env["views"][2]
    a = tb.render(RES, RES, spp=1, camera_matrix=v["xform"], focal_length=v["focal_length"])
    b = twin.render(RES, RES, spp=1, camera_matrix=v["xform"], focal_length=v["focal_length"])
    assert float(np.mean((a[..., :3] - b[..., :3]) ** 2)) < 0.01                 # both learned the same scene
    assert tb.nerf.training.n_steps_between_error_map_updates > 100          # 16 -> 24 -> 36 -> ... : the updates thin out
    assert int(tb.get_option("error_cdf_width")) > 0
    tb.nerf.training.sample_focal_plane_proportional_to_error = False
    tb.nerf.training.sample_image_proportional_to_error = False
    for _ in range(40):
        tb.train(BATCH)                                                       # uniform sampling again (prefetched rays of the old mode are discarded)
    assert np.isfinite(tb.loss)


def test_camera_gradients_are_divided_by_the_pixel_pdf(env, ref):
    """compute_cam_gradient_train_nerf with the focal-plane CDFs in use (testbed_nerf.cu:2083-2107): every ray's contribution / uv_pdf"""
    from .test_extrinsics_gpu import _step_split
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tr = tb.nerf.training
    tr.sample_focal_plane_proportional_to_error = True
    tr.n_steps_between_error_map_updates = 4
    for _ in range(4):
        tb.train(BATCH)
    b = _buffers(tb)
    tr.optimize_extrinsics = True
    tr.n_steps_between_cam_updates = 1
    st = tb.stats
    rpb, step = st.rays_per_batch, st.training_step
    finish = _step_split(tb)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    nr = int(c[0])
    ray_indices = tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32).copy()
    rays = tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6).copy()
    numsteps = tb.debug_tensor("numsteps", torch.int32, nr * 2).cpu().numpy().view(np.uint32).reshape(nr, 2).copy()
    coords = tb.debug_tensor("coords_compacted", torch.float32, BATCH * 7).cpu().numpy().reshape(BATCH, 7).copy()
    cgrad = tb.debug_tensor("coords_gradient", torch.float32, BATCH * 7).cpu().numpy().reshape(BATCH, 7).copy()
    gp = tb.debug_tensor("cam_pos_gradient", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3).copy()
    gr = tb.debug_tensor("cam_rot_gradient", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3).copy()
    finish()
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    plain_p, plain_r = ref.compute_cam_gradient(rpb, aabb, N_VIEWS, nr, ray_indices, rays, numsteps, coords, cgrad)
    ref.set_error_map(b["cdf_x_cond_y"], b["cdf_y"], None, b["cdf_res"])
    ref.set_cam_gradient_rng(1337, 1 + step * (1 << 32), env["images"], True)
    try:
        wp, wr = ref.compute_cam_gradient(rpb, aabb, N_VIEWS, nr, ray_indices, rays, numsteps, coords, cgrad)
    finally:
        ref.set_error_map(); ref.set_cam_gradient_rng(0, 0, None, 0)
    assert np.abs(gp - wp).max() <= 2e-3 * np.abs(wp).max() and np.abs(gr - wr).max() <= 2e-3 * np.abs(wr).max()
    assert np.abs(plain_p - wp).max() > 20 * max(np.abs(gp - wp).max(), 1e-12)                 # the division is there
