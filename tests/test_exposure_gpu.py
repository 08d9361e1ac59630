"""nerf.training.optimize_exposure (the exposure part): the loss kernel's per-image exposure scale and its gradient
(testbed_nerf.cu:1794,1801-1818,1969-1983) against the oracle on the GPU's own network output, the device-side AdamOptimizer<vec3> step with
mean renormalisation (adam_optimizer.h:143-150, testbed_nerf.cu:3831-3858) against a numpy restatement over a trajectory, and an end-to-end
run that recovers per-image brightness offsets baked into the training views.  Also the colour-space branch of the training target
(Testbed::m_color_space: Linear by default, :1801-1808; SRGB :1809-1818) on views with semi-transparent pixels."""
import importlib

import numpy as np
import pytest
import torch

from .test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG, _step_and_capture

pytestmark = pytest.mark.gpu

LR = 1e-2          # SMALL_CFG's Adam learning rate; the exposure optimisers follow the main optimizer's current rate (:3843)


@pytest.fixture(scope="module")
def env(ngp, ref):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    images, keep = ref.images_from_views(views)
    return dict(pyngp=pyngp, scene=scene, views=views, images=images, keep=keep)


def _oracle_loss(ref, env, tb, rpb, max_inf, step, exposure, color_space=0, linear_colors=0):
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    bf = tb.density_grid(1).cpu().numpy()
    skip = 1 + step * (1 << 32)
    o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
    nr = int(o["ray_counter"]); n_written = int(o["numsteps"][:nr, 0].sum())
    mlp_out = tb.debug_tensor("mlp_out", torch.float16, n_written * 4).cpu().numpy().reshape(n_written, 4)
    mean_density = float(tb.debug_tensor("mean_density", torch.float32, 1).cpu().numpy()[0])
    cfg = ref.LossCfg(4, 2, 3, 1, linear_colors, 1, (ref.f32 * 3)(0, 0, 0), 128.0, 0.1, mean_density)
    cfg.color_space = color_space
    grad = np.zeros((N_VIEWS, 3), np.float32)
    exposure = None if exposure is None else np.ascontiguousarray(exposure, np.float32)
    if exposure is not None:
        cfg.exposure = ref.P(exposure).value; cfg.exposure_gradient = ref.P(grad).value
    lo = ref.compute_loss(rpb, aabb, 1337, skip, BATCH, nr, cfg, env["images"], mlp_out, o["ray_indices"], o["rays"], o["numsteps"], o["coords"])
    return lo, grad, nr


def _check_dloss(tb, lo, nr):
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    assert c[2] == lo["compacted_counter"]
    nc = int(min(c[2], BATCH))
    gdl = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    scale = np.abs(rdl).max()
    assert np.abs(gdl[:nc] - rdl).max() <= 2e-3 * scale + 1e-7, (np.abs(gdl[:nc] - rdl).max(), scale)
    return nc


@pytest.mark.parametrize("lanes", [1, 8, 32])
def test_exposure_gradient_and_adam_trajectory(env, ref, lanes):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.nerf.training.optimize_exposure = True
    tb.nerf.training.n_steps_between_cam_updates = 1            # an optimiser step after every training step: the gradient buffer holds one step
    assert tb.nerf.training.optimize_exposure is True and tb.nerf.training.exposure_l2_reg == 0.0
    e = np.zeros((N_VIEWS, 3), np.float32); m1 = np.zeros_like(e); m2 = np.zeros_like(e)
    for it in range(1, 7):
        rpb, max_inf, step = _step_and_capture(tb, ref, env, 0)
        g_gpu = tb.debug_tensor("cam_exposure_gradient", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3).copy()
        # --- loss kernel: targets scaled by 2^e, gradient of the exposures, against the oracle fed with the exposures this step used ---
        lo, g_ref, nr = _oracle_loss(ref, env, tb, rpb, max_inf, step, e)
        _check_dloss(tb, lo, nr)
        scale = np.abs(g_ref).max()
        assert scale > 0
        assert np.abs(g_gpu - g_ref).max() <= 2e-3 * scale, (it, np.abs(g_gpu - g_ref).max(), scale)      # float atomics in another order + fp32 re-association
        # --- AdamOptimizer<vec3>::step per image on gradient * per_camera_loss_scale, then minus the mean (numpy float32 restatement) ---
        g = g_gpu * np.float32(N_VIEWS / 128.0 / 1.0)
        lr = np.float32(LR) * np.sqrt(np.float32(1.0) - np.float32(0.99) ** np.float32(it)) / (np.float32(1.0) - np.float32(0.9) ** np.float32(it))
        m1 = np.float32(0.9) * m1 + np.float32(1.0 - 0.9) * g
        m2 = np.float32(0.99) * m2 + np.float32(1.0 - 0.99) * g * g
        e = e - np.float32(lr) * m1 / (np.sqrt(m2) + np.float32(1e-8))
        e = (e - e.mean(axis=0, keepdims=True)).astype(np.float32)
        e_gpu = tb.debug_tensor("cam_exposure", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3)
        np.testing.assert_allclose(e_gpu, e, rtol=2e-4, atol=2e-6)
        assert abs(float(e_gpu.mean())) < 1e-6
        e = e_gpu.copy()                                         # follow the device's values: the next step's oracle must see what the kernel saw
    assert np.abs(e).max() > 1e-3                                # the exposures did move


def test_exposure_is_off_by_default_and_costs_nothing(env, ref):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    assert tb.nerf.training.optimize_exposure is False and tb.nerf.training.n_steps_between_cam_updates == 16
    rpb, max_inf, step = _step_and_capture(tb, ref, env, 3)
    ptr, n = tb.debug_buffer("cam_exposure")
    assert ptr is None or ptr == 0                               # never allocated
    lo, _, nr = _oracle_loss(ref, env, tb, rpb, max_inf, step, None)
    _check_dloss(tb, lo, nr)


@pytest.fixture(scope="module")
def env_translucent(env, ref):
    """the same views with every alpha halved: the three target formulas differ wherever 0 < alpha < 1"""
    views = []
    for v in env["views"]:
        px = v["pixels"].clone()
        px[..., 3] = px[..., 3] // 2
        views.append({**v, "pixels": px})
    images, keep = ref.images_from_views(views)
    return dict(env, views=views, images=images, keep=keep)


@pytest.mark.parametrize("color_space,linear_colors", [(0, 0), (1, 0), (0, 1)])
def test_training_target_colour_space_branches(env_translucent, ref, color_space, linear_colors):
    """the three target formulas (:1801-1818): Linear colour space (the default), SRGB colour space, training in linear colours"""
    env = env_translucent
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.color_space = color_space
    tb.nerf.training.linear_colors = bool(linear_colors)
    rpb, max_inf, step = _step_and_capture(tb, ref, env, 3)
    lo, _, nr = _oracle_loss(ref, env, tb, rpb, max_inf, step, None, color_space, linear_colors)
    _check_dloss(tb, lo, nr)
    gl = tb.debug_tensor("loss", torch.float32, rpb).cpu().numpy()
    np.testing.assert_allclose(gl[:nr], lo["loss"][:nr], rtol=1e-3, atol=1e-9)
    # the formula matters on these views: the other colour space gives other losses
    other, _, _ = _oracle_loss(ref, env, tb, rpb, max_inf, step, None, 1 - color_space, 0)
    assert (np.abs(other["loss"][:nr] - lo["loss"][:nr]) > 1e-3 * np.abs(lo["loss"][:nr])).mean() > 0.01      # the rays that hit a translucent pixel


def test_exposure_optimisation_recovers_per_image_brightness(env):
    """Every camera appears twice: once darkened by one stop and once brightened by one (in linear light).  The network sees the same rays
    for both copies and cannot tell them apart, so only the per-image exposures can explain the difference: with optimize_exposure they
    converge to +1 / -1 stop (the renormalisation pins their mean to 0).  The sign and most of the magnitude must be there."""
    pyngp, scene = env["pyngp"], env["scene"]
    n_cam, res = 6, 64
    base = scene.make_views(n_cam, res, res)
    views, stops = [], []
    for v in base:
        px = v["pixels"].numpy().astype(np.float32) / 255.0
        lin = np.where(px[..., :3] <= 0.04045, px[..., :3] / 12.92, ((px[..., :3] + 0.055) / 1.055) ** 2.4)
        for e in (1.0, -1.0):                                   # the exposure the optimiser should find for this copy
            scaled = lin * (2.0 ** -e)
            srgb = np.where(scaled < 0.0031308, 12.92 * scaled, 1.055 * np.power(np.maximum(scaled, 1e-12), 1 / 2.4) - 0.055)
            q = v["pixels"].clone()
            q.numpy()[..., :3] = np.clip(np.round(srgb * 255.0), 0, 255).astype(np.uint8)
            views.append({**v, "pixels": q}); stops.append(e)
    stops = np.array(stops, np.float32); n = len(views)
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(views, aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.nerf.training.optimize_exposure = True
    tb.nerf.training.n_steps_between_cam_updates = 4            # an Adam step moves an exposure by at most ~lr = 0.01 stop: 500 updates are ample for 1 stop
    for _ in range(2000):
        tb.train(1 << 14)
    e = tb.debug_tensor("cam_exposure", torch.float32, n * 3).cpu().numpy().reshape(n, 3).mean(axis=1)
    print("learned exposures (stops):", np.round(e, 3))
    assert np.all(np.sign(e) == np.sign(stops)), e
    assert np.abs(e).min() > 0.5 and np.abs(e).max() < 1.5, e
    assert abs(float(e.mean())) < 1e-3
