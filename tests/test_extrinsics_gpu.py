"""nerf.training.optimize_extrinsics (the camera part): the backward pass returns dL/d(coords) (NerfNetwork::backward's
dL_dinput), compute_cam_gradient_train_nerf (testbed_nerf.cu:2014-2123) sums it per ray into dL/d(origin) and dL/d(direction) and per view into a
position and an angle-axis gradient, and every n_steps_between_cam_updates steps the host takes one Adam step per view on a position offset
(AdamOptimizer<vec3>) and a rotation offset (RotationAdamOptimizer), after which update_transforms (:2984-3023) re-poses the views.
Each stage against the oracle / a numpy restatement, then an end-to-end run that pulls perturbed cameras back."""
import importlib

import numpy as np
import pytest
import torch

from .test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG

pytestmark = pytest.mark.gpu

LR = 1e-2          # SMALL_CFG's Adam learning rate (the floor of the extrinsic learning rate is a thousandth of it)


@pytest.fixture(scope="module")
def env(ngp, ref):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    return dict(pyngp=pyngp, scene=scene, views=views)


def _rotmat(v):
    """Rodrigues' formula in float64 (only used to build test poses; the parity checks go through the oracle's pinned rotmat / rotvec)"""
    v = np.asarray(v, np.float64)
    angle = np.linalg.norm(v)
    if angle == 0:
        return np.eye(3)
    k = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]]) / angle
    return np.eye(3) + np.sin(angle) * k + (1 - np.cos(angle)) * (k @ k)


def _mul3(A, B):
    """glm mat3 * mat3 in float32, products summed left to right"""
    f = np.float32
    out = np.zeros((3, B.shape[1]), np.float32)
    for c in range(B.shape[1]):
        for r in range(3):
            out[r, c] = f(f(f(A[r, 0] * B[0, c]) + f(A[r, 1] * B[1, c])) + f(A[r, 2] * B[2, c]))
    return out


def _step_split(tb):
    """one training step in its two halves: the parameter / camera gradients exist between them"""
    import ctypes as C
    _lib = importlib.import_module("instant-ngp_b200._lib")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().ngp_testbed_train_begin(tb._h, stream, BATCH))
    torch.cuda.synchronize()
    m, mb = C.c_uint32(), C.c_uint32()
    _lib.check(_lib.lib().ngp_testbed_pending_counters(tb._h, C.byref(m), C.byref(mb)))
    return lambda: _lib.check(_lib.lib().ngp_testbed_train_end(tb._h, stream, max(m.value, 1), max(mb.value, 1)))


@pytest.mark.parametrize("lanes,n_warm", [(1, 3), (8, 5), (0, 40)])
def test_camera_gradients_of_one_step(env, ref, lanes, n_warm):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    assert tb.nerf.training.optimize_extrinsics is False
    tb.nerf.training.optimize_extrinsics = True
    tb.nerf.training.n_steps_between_cam_updates = 1          # every step is a window of its own: the accumulators hold one step
    for _ in range(n_warm):
        tb.train(BATCH)
    params = tb.param_buffer(1).cpu().numpy().copy()
    rpb = tb.stats.rays_per_batch
    finish = _step_split(tb)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    nr = int(c[0]); nc = int(min(c[2], BATCH))
    ray_indices = tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32).copy()
    rays = tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6).copy()
    numsteps = tb.debug_tensor("numsteps", torch.int32, nr * 2).cpu().numpy().view(np.uint32).reshape(nr, 2).copy()
    coords = tb.debug_tensor("coords_compacted", torch.float32, BATCH * 7).cpu().numpy().reshape(BATCH, 7).copy()
    dloss = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).copy()
    cgrad = tb.debug_tensor("coords_gradient", torch.float32, BATCH * 7).cpu().numpy().reshape(BATCH, 7).copy()
    gp = tb.debug_tensor("cam_pos_gradient", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3).copy()
    gr = tb.debug_tensor("cam_rot_gradient", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3).copy()
    finish()

    # dL/d(coords) of the step: the oracle network on a slice of the compacted batch
    rcfg = ref.nerf_cfg_from_json(SMALL_CFG)
    k = min(nc, 2048)
    _, want = ref.nerf_backward_input(rcfg, params, coords[:k], dloss[:k])
    for cols in (slice(0, 3), slice(4, 7)):
        a, b = cgrad[:k, cols], want[:, cols]
        assert np.linalg.norm(a - b) <= 2e-2 * np.linalg.norm(b) + 1e-12, (cols, np.linalg.norm(a - b), np.linalg.norm(b))
    # per-view sums: the oracle kernel on the device's own per-sample gradients (float atomics in another order)
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    wp, wr = ref.compute_cam_gradient(rpb, aabb, N_VIEWS, nr, ray_indices, rays, numsteps, coords, cgrad)
    assert np.abs(wp).max() > 0 and np.abs(wr).max() > 0
    assert np.abs(gp - wp).max() <= 2e-3 * np.abs(wp).max(), (np.abs(gp - wp).max(), np.abs(wp).max())
    assert np.abs(gr - wr).max() <= 2e-3 * np.abs(wr).max(), (np.abs(gr - wr).max(), np.abs(wr).max())
    # views whose rays did not fit into the batch get nothing
    img_of_ray = (np.arange(rpb, dtype=np.int64) * N_VIEWS // rpb)[ray_indices]
    seen = np.unique(img_of_ray[numsteps[:, 0] > 0])
    unseen = np.setdiff1d(np.arange(N_VIEWS), seen)
    assert np.all(gp[unseen] == 0) and np.all(gr[unseen] == 0)


def test_pose_optimisers_follow_the_reference_arithmetic(env, ref):
    """AdamOptimizer<vec3> on the position offset, RotationAdamOptimizer on the rotation offset (adam_optimizer.h:143-150,238-247), the decaying
    learning rate and the l2 term (testbed_nerf.cu:3790-3804), update_transforms (:3011-3021): numpy restatement fed with the device's gradients"""
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.nerf.training.optimize_extrinsics = True
    tb.nerf.training.n_steps_between_cam_updates = 1
    assert abs(tb.nerf.training.extrinsic_l2_reg - 1e-4) < 1e-10 and abs(tb.nerf.training.extrinsic_learning_rate - 1e-3) < 1e-9
    base = [np.asarray(v["xform"], np.float32).reshape(4, 3).T.copy() for v in env["views"]]     # 3 x 4, NGP convention
    for i in range(N_VIEWS):
        np.testing.assert_array_equal(tb.get_camera_extrinsics(i, convert_to_ngp=False), base[i])
    pos = np.zeros((N_VIEWS, 3), np.float32); rot = np.zeros((N_VIEWS, 3), np.float32)
    m1p = np.zeros_like(pos); m2p = np.zeros_like(pos); m1r = np.zeros_like(pos); m2r = np.zeros_like(pos)
    f = np.float32
    for it in range(1, 7):
        finish = _step_split(tb)
        gp = tb.debug_tensor("cam_pos_gradient", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3).copy()
        gr = tb.debug_tensor("cam_rot_gradient", torch.float32, N_VIEWS * 3).cpu().numpy().reshape(N_VIEWS, 3).copy()
        finish()
        scale = f(N_VIEWS) / f(128.0) / f(1.0)
        lr = max(f(1e-3) * f(0.33) ** f((it - 1) // 128), f(LR) / f(1000.0))
        actual = f(lr) * np.sqrt(f(1) - f(0.99) ** f(it)) / (f(1) - f(0.9) ** f(it))
        g = gp * scale + pos * f(1e-4)
        m1p = f(0.9) * m1p + f(0.1) * g; m2p = f(0.99) * m2p + f(0.01) * g * g
        pos = (pos - actual * m1p / (np.sqrt(m2p) + f(1e-8))).astype(np.float32)
        g = gr * scale + rot * f(1e-4)
        m1r = f(0.9) * m1r + f(0.1) * g; m2r = f(0.99) * m2r + f(0.01) * g * g
        step = (actual * m1r / (np.sqrt(m2r) + f(1e-8))).astype(np.float32)
        rot = np.stack([ref.rotvec(_mul3(ref.rotmat(-step[i]), ref.rotmat(rot[i]))) for i in range(N_VIEWS)])   # pinned against the reference's GLM
        for i in range(N_VIEWS):
            want = np.concatenate([_mul3(ref.rotmat(rot[i]), base[i][:, :3]), (base[i][:, 3] + pos[i])[:, None]], axis=1)
            got = tb.get_camera_extrinsics(i, convert_to_ngp=False)
            np.testing.assert_allclose(got, want, rtol=0, atol=3e-6, err_msg=f"view {i} after update {it}")
    assert np.abs(pos).max() > 1e-4 and np.abs(rot).max() > 1e-4                 # the cameras did move
    # the NeRF-convention read-out inverts nerf_matrix_to_ngp (scale 0.33, offset 0.5 by default)
    m = tb.get_camera_extrinsics(0)
    ngp = tb.get_camera_extrinsics(0, convert_to_ngp=False)
    back = np.stack([ngp[[2, 0, 1], 0], -ngp[[2, 0, 1], 1], -ngp[[2, 0, 1], 2], (ngp[[2, 0, 1], 3] - 0.5) / 0.33], axis=1)   # ngp_matrix_to_nerf (nerf_loader.h:142-158)
    np.testing.assert_allclose(m, back, atol=1e-6)
    # reset: the given poses are in force again
    tb.reset()
    for i in range(N_VIEWS):
        np.testing.assert_array_equal(tb.get_camera_extrinsics(i, convert_to_ngp=False), base[i])


def _pose_error(tb, truth):
    err_r, err_t = [], []
    for i, t in enumerate(truth):
        m = tb.get_camera_extrinsics(i, convert_to_ngp=False).astype(np.float64)
        err_t.append(np.linalg.norm(m[:, 3] - t[:, 3]))
        err_r.append(np.degrees(np.arccos(np.clip((np.trace(m[:, :3] @ t[:, :3].T) - 1) / 2, -1, 1))))
    return float(np.mean(err_r)), float(np.mean(err_t))


def test_perturbed_cameras_are_pulled_back(env):
    """Half of the cameras are rotated by 1.5 degrees and shifted by 0.015 (scene units); trained with optimize_extrinsics the mean pose error
    of the perturbed views falls, while the untouched ones stay close to where they were."""
    pyngp, scene = env["pyngp"], env["scene"]
    n, res = 16, 96
    views = scene.make_views(n, res, res)
    truth = [np.asarray(v["xform"], np.float64).reshape(4, 3).T.copy() for v in views]
    rng = np.random.default_rng(4)
    perturbed = []
    for i, v in enumerate(views):
        m = truth[i].copy()
        if i % 2 == 0:
            axis = rng.normal(size=3); axis /= np.linalg.norm(axis)
            m[:, :3] = _rotmat(axis * np.radians(1.5)) @ m[:, :3]
            m[:, 3] += rng.normal(size=3) / np.sqrt(3) * 0.015
        perturbed.append({**v, "xform": np.ascontiguousarray(m.T.reshape(-1), np.float32)})
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(perturbed, aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.nerf.training.optimize_extrinsics = True
    even = list(range(0, n, 2)); odd = list(range(1, n, 2))
    def errors():
        r, t = zip(*[_pose_error_one(tb, i, truth[i]) for i in range(n)])
        return np.array(r), np.array(t)
    def _pose_error_one(tb, i, t):
        m = tb.get_camera_extrinsics(i, convert_to_ngp=False).astype(np.float64)
        return (np.degrees(np.arccos(np.clip((np.trace(m[:, :3] @ t[:, :3].T) - 1) / 2, -1, 1))), np.linalg.norm(m[:, 3] - t[:, 3]))
    r0, t0 = errors()
    for _ in range(3000):
        tb.train(1 << 14)
    r1, t1 = errors()
    print("rotation error (deg) before / after, perturbed views:", r0[even].mean(), r1[even].mean(), " untouched:", r0[odd].mean(), r1[odd].mean())
    print("position error before / after, perturbed views:", t0[even].mean(), t1[even].mean(), " untouched:", t0[odd].mean(), t1[odd].mean())
    # (the scene fixes the cameras only relative to each other: the untouched half gives a little while the perturbed half comes back)
    assert r1[even].mean() < 0.7 * r0[even].mean() and t1[even].mean() < 0.85 * t0[even].mean()
    assert r1[odd].mean() < 0.5 and t1[odd].mean() < 0.01
