"""nerf.training.optimize_distortion (the distortion buffer): a lattice of 2-D image-plane offsets shared by all views
(TrainableBuffer<2, 2, float>, config "distortion_map"; testbed.cu:4066-4077) is added to every training ray's direction (common_device.cuh:491-493),
receives the rays' direction gradients projected into the image plane (testbed_nerf.cu:2088-2100, deposit_image_gradient) and is stepped by its own
Adam at the camera-update cadence (:3811-3818).  Rays under a non-zero map, the gradient splats and the optimiser against the oracle / numpy."""
import importlib

import numpy as np
import pytest
import torch

from .test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG
from .test_extrinsics_gpu import _step_split

pytestmark = pytest.mark.gpu

DW, DH = 8, 6
CFG = {**SMALL_CFG, "distortion_map": {"resolution": [DW, DH], "optimizer": {"otype": "ExponentialDecay", "decay_start": 2, "decay_interval": 2, "decay_base": 0.5,
                                                                              "nested": {"otype": "Adam", "learning_rate": 2e-3, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-8}}}}


@pytest.fixture(scope="module")
def env(ngp, ref):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    images, keep = ref.images_from_views(views)
    return dict(pyngp=pyngp, scene=scene, views=views, images=images, keep=keep)


def _map(tb):
    return tb.debug_tensor("distortion_map", torch.float32, DW * DH * 2).cpu().numpy().reshape(DH, DW, 2).copy()


@pytest.mark.parametrize("lanes", [1, 8, 0])
def test_distortion_map_rays_gradients_and_optimiser(env, ref, lanes):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(CFG)
    tr = tb.nerf.training
    assert tr.optimize_distortion is False
    for _ in range(3):
        tb.train(BATCH)
    tr.optimize_distortion = True
    tr.n_steps_between_cam_updates = 1
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    f = np.float32
    m1 = np.zeros((DH, DW, 2), np.float32); m2 = np.zeros_like(m1); steps = np.zeros((DH, DW, 2), np.int64)
    want_map = np.zeros((DH, DW, 2), np.float32)
    lr = f(2e-3)
    for it in range(1, 6):
        st = tb.stats
        rpb, mb, step = st.rays_per_batch, st.measured_batch_size_before_compaction, st.training_step
        max_inf = BATCH * 16 if mb == 0 else (min(mb, BATCH * 16) + 255) // 256 * 256
        finish = _step_split(tb)
        before = _map(tb)                                            # the map this step's rays went through
        np.testing.assert_allclose(before, want_map, rtol=1e-5, atol=1e-9)
        c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
        nr = int(c[0])
        ray_indices = tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32).copy()
        rays = tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6).copy()
        numsteps = tb.debug_tensor("numsteps", torch.int32, nr * 2).cpu().numpy().view(np.uint32).reshape(nr, 2).copy()
        coords = tb.debug_tensor("coords_compacted", torch.float32, BATCH * 7).cpu().numpy().reshape(BATCH, 7).copy()
        cgrad = tb.debug_tensor("coords_gradient", torch.float32, BATCH * 7).cpu().numpy().reshape(BATCH, 7).copy()
        g_dev = tb.debug_tensor("distortion_gradient", torch.float32, DW * DH * 2).cpu().numpy().reshape(DH, DW, 2).copy()
        w_dev = tb.debug_tensor("distortion_gradient_weight", torch.float32, DW * DH * 2).cpu().numpy().reshape(DH, DW, 2).copy()
        bf = tb.density_grid(1).cpu().numpy()
        finish()
        skip = 1 + step * (1 << 32)
        # --- rays of the step: through the map as it stood ---
        g_ref = np.zeros((DH, DW, 2), np.float32); w_ref = np.zeros((DH, DW, 2), np.float32)
        ref.set_distortion(before, g_ref, w_ref)
        ref.set_cam_gradient_rng(1337, skip, env["images"], True)
        try:
            o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
            assert int(o["ray_counter"]) == nr
            assert np.array_equal(rays, o["rays"][:nr])
            # --- gradient and weight splats of the step (float atomics in another order) ---
            ref.compute_cam_gradient(rpb, aabb, N_VIEWS, nr, ray_indices, rays, numsteps, coords, cgrad)
        finally:
            ref.set_distortion(); ref.set_cam_gradient_rng(0, 0, None, 0)
        if it > 1:
            plain = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
            k = min(nr, int(plain["ray_counter"]))
            assert not np.array_equal(plain["rays"][:k, 3:], rays[:k, 3:])           # a non-zero map bends the rays
        assert w_ref.sum() > 0 and np.abs(g_ref).max() > 0
        np.testing.assert_allclose(w_dev, w_ref, rtol=1e-4, atol=1e-5)
        assert np.abs(g_dev - g_ref).max() <= 2e-3 * np.abs(g_ref).max()
        # --- safe_divide + Adam (non-matrix parameters: zero gradients are skipped, no L2), loss scale 128 x n_steps_between_cam_updates ---
        with np.errstate(all="ignore"):
            g = np.where(w_dev > 0, g_dev / w_dev, f(0)).astype(np.float32) / f(128.0)
        live = g != 0
        steps[live] += 1
        s = np.maximum(steps, 1).astype(np.float32)
        lr_t = (lr * np.sqrt(f(1) - f(0.99) ** s) / (f(1) - f(0.9) ** s)).astype(np.float32)
        m1 = np.where(live, f(0.9) * m1 + f(0.1) * g, m1).astype(np.float32)
        m2 = np.where(live, f(0.99) * m2 + f(0.01) * g * g, m2).astype(np.float32)
        want_map = np.where(live, want_map - lr_t / (np.sqrt(m2) + f(1e-8)) * m1, want_map).astype(np.float32)
        if it >= 2 and (it - 2) % 2 == 0:                                            # ExponentialDecay: x 0.5 at optimiser steps 2, 4, ...
            lr = f(lr * f(0.5))
    np.testing.assert_allclose(_map(tb), want_map, rtol=1e-4, atol=1e-8)
    assert np.abs(want_map).max() > 1e-4
    # the tracer bends its rays by the trained map only with render_with_lens_distortion (testbed_nerf.cu:2776-2779)
    tb.set_camera_to_training_view(1)                                # switches render_with_lens_distortion on (testbed.cu:855)
    assert tb.nerf.render_with_lens_distortion is True
    b = tb.render(RES, RES, 1, True)
    tb.nerf.render_with_lens_distortion = False
    a = tb.render(RES, RES, 1, True)
    assert np.isfinite(b).all() and not np.array_equal(a, b) and np.abs(a - b).mean() < 0.05
    # turning the optimisation off keeps the trained map in the rays; a reset clears it
    tr.optimize_distortion = False
    tb.train(BATCH)
    assert np.array_equal(_map(tb), _map(tb))
    tb.reset()
    tb.train(BATCH)
