"""GPU parity of the NeRF training step / occupancy grid / tracer through the Testbed C ABI vs the CPU oracle.
Integer paths (grid sample cells, bitfield, ray->sample compaction, traversal with cone_angle 0) are compared bit-exact."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SMALL_CFG = {
    "loss": {"otype": "Huber"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.3819},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
BATCH = 1 << 14
N_VIEWS, RES = 12, 96


@pytest.fixture(scope="module")
def env(ngp, ref):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    images, keep = ref.images_from_views(views)
    return dict(pyngp=pyngp, scene=scene, views=views, images=images, keep=keep)


def _testbed(env, lanes=0, prefetch=True, cfg=None):
    tb = env["pyngp"].Testbed(env["pyngp"].TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes     # 0 automatic; 1 / 2 / 4 / 8 / 32 lanes cooperate on one ray (identical results)
    tb.prefetch_rays = prefetch          # next step's rays generated beside this step's backward pass (identical results)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(cfg or SMALL_CFG)
    return tb


def test_synthetic_views_match_cpu_generator(env, ref):
    scene = env["scene"]
    cpu = ref.synth_views(N_VIEWS, RES, RES, scene.RADIUS, scene.focal_px(RES), step=1.73205080757 / 4096.0, count=2)
    for i in range(2):
        a = env["views"][i]["pixels"].numpy().astype(np.int32); b = cpu[i]["pixels"].astype(np.int32)
        assert np.abs(a[..., 3] - b[..., 3]).max() <= 2                      # alpha: expf last-ulp differences only
        solid = (a[..., 3] > 32) & (b[..., 3] > 32)                          # straight colour is ill-conditioned where alpha ~ 0
        assert np.abs(a[solid][:, :3] - b[solid][:, :3]).max() <= 3
        np.testing.assert_allclose(env["views"][i]["xform"], cpu[i]["xform"], atol=1e-6)
    assert (env["views"][0]["pixels"].numpy()[..., 3] > 0).mean() > 0.1


def test_occupancy_grid_first_update(env, ref):
    tb = _testbed(env)
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    params0 = tb.param_buffer(1).cpu().numpy().copy()
    tb.train(BATCH)
    NC = 128 ** 3
    # grid sample cells + positions of the step-0 update: bit-exact (pcg32 + LCG cell pick + morton)
    grid0 = ref.mark_untrained(np.zeros(NC, np.float32), env["images"])
    dg_seed = np.zeros(1, np.uint32); ref.lib().ref_pcg32_fill_uint(ref.u64(1337), ref.u64(0), ref.P(dg_seed), ref.C.c_size_t(1))
    pos, idx = ref.generate_grid_samples(NC, int(dg_seed[0]), 0, 0, aabb, grid0, 1, -0.01)
    gidx = tb.debug_tensor("grid_idx", torch.int32, NC).cpu().numpy().view(np.uint32)
    gpos = tb.debug_tensor("grid_pos", torch.float32, NC * 3).cpu().numpy().reshape(NC, 3)
    assert np.array_equal(gidx, idx)
    assert np.array_equal(gpos, pos)
    # densities of the first 4096 samples vs the oracle network on the initial parameters
    rcfg = ref.nerf_cfg_from_json(SMALL_CFG)
    dens = tb.debug_tensor("grid_out", torch.float16, NC).cpu().numpy()[:4096].astype(np.float32)
    dref = ref.nerf_density(rcfg, params0, pos[:4096].copy()).astype(np.float32)
    np.testing.assert_allclose(dens, dref, rtol=2e-3, atol=2e-3)
    # bitfield + mips: bit-exact given the GPU's density grid
    grid = tb.density_grid(0).cpu().numpy()
    bf, mean = ref.update_bitfield(grid, 1)
    assert np.array_equal(tb.density_grid(1).cpu().numpy(), bf)
    gmean = tb.debug_tensor("mean_density", torch.float32, 1).cpu().numpy()[0]
    assert abs(gmean - mean) <= 1e-6 * max(1.0, abs(mean))
    # cells no camera sees stay -1, visible ones hold max(1*decay, density)
    assert ((grid == -1) == (grid0 == -1)).all()


def _step_and_capture(tb, ref, env, n_warm, batch=BATCH):
    for _ in range(n_warm):
        tb.train(batch)
    st = tb.stats
    rpb, mb = st.rays_per_batch, st.measured_batch_size_before_compaction
    max_samples = batch * 16
    max_inference = max_samples if mb == 0 else (min(mb, max_samples) + 255) // 256 * 256
    step = st.training_step
    tb.train(batch)
    return rpb, max_inference, step


def _check_step_against_oracle(tb, ref, env, n_warm, prefetch, batch):
    """one captured training step: rays, per-ray counts, every sample coordinate, compacted counts and coordinates bit-exact;
    loss gradients within 2e-3 of their scale; fill_rollover semantics"""
    rpb, max_inf, step = _step_and_capture(tb, ref, env, n_warm, batch)
    if n_warm > 32:
        assert (tb.prefetch_hits > 0) == prefetch
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    bf = tb.density_grid(1).cpu().numpy()
    skip = 1 + step * (1 << 32)
    o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    assert c[0] == o["ray_counter"] and c[1] == o["numsteps_counter"]
    nr = int(c[0]); ns = int(min(c[1], max_inf))
    assert nr > 0 and ns > 0
    assert np.array_equal(tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32), o["ray_indices"][:nr])
    assert np.array_equal(tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6), o["rays"][:nr])
    n_written = int(o["numsteps"][:nr, 0].sum())
    gcoords = tb.debug_tensor("coords", torch.float32, n_written * 7).cpu().numpy().reshape(n_written, 7)
    assert np.array_equal(gcoords, o["coords"][:n_written])          # traversal bit-exact (cone_angle 0)

    # loss / compaction on the GPU's own network output
    mlp_out = tb.debug_tensor("mlp_out", torch.float16, n_written * 4).cpu().numpy().reshape(n_written, 4)
    mean_density = float(tb.debug_tensor("mean_density", torch.float32, 1).cpu().numpy()[0])
    cfg = ref.LossCfg(4, 2, 3, 1, 0, 1, (ref.f32 * 3)(0, 0, 0), 128.0, 0.1, mean_density)
    lo = ref.compute_loss(rpb, aabb, 1337, skip, batch, nr, cfg, env["images"], mlp_out, o["ray_indices"], o["rays"], o["numsteps"], o["coords"])
    assert c[2] == lo["compacted_counter"]
    gns = tb.debug_tensor("numsteps", torch.int32, nr * 2).cpu().numpy().view(np.uint32).reshape(nr, 2)
    assert np.array_equal(gns, lo["numsteps"][:nr])
    nc = int(min(c[2], batch))
    gcc = tb.debug_tensor("coords_compacted", torch.float32, batch * 7).cpu().numpy().reshape(batch, 7)
    assert np.array_equal(gcc[:nc], lo["coords_compacted"][:nc])
    gdl = tb.debug_tensor("dloss", torch.float16, batch * 4).cpu().numpy().reshape(batch, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    scale = np.abs(rdl).max()
    assert np.abs(gdl[:nc] - rdl).max() <= 2e-3 * scale + 1e-7, (np.abs(gdl[:nc] - rdl).max(), scale)
    gl = tb.debug_tensor("loss", torch.float32, rpb).cpu().numpy()
    np.testing.assert_allclose(gl[:nr], lo["loss"][:nr], rtol=1e-3, atol=1e-9)
    # fill_rollover: rows [nc, batch) are cyclic copies, gradients scaled by nc/batch
    if nc < batch:
        src = np.arange(nc, batch) % nc
        assert np.array_equal(gcc[nc:], gcc[src])
        np.testing.assert_allclose(gdl[nc:], gdl[src] * (nc / batch), rtol=2e-3, atol=1e-7)
    return nr, n_written, nc


@pytest.mark.parametrize("n_warm,lanes,prefetch", [(w, l, True) for w in (0, 5) for l in (1, 2, 4, 8, 32)] + [(37, 0, True), (37, 0, False), (51, 8, True)])
def test_sampler_and_loss_bit_exact_compaction(env, ref, n_warm, lanes, prefetch):
    """n_warm >= 33 and odd: the captured step does not refresh the occupancy grid, so with prefetch its rays were generated
    on the second stream during the step before"""
    tb = _testbed(env, lanes, prefetch)
    _check_step_against_oracle(tb, ref, env, n_warm, prefetch, BATCH)


HEADLINE_CFG = {**SMALL_CFG, "encoding": {**SMALL_CFG["encoding"], "log2_hashmap_size": 19}}


@pytest.mark.parametrize("n_warm", [2, 37])
def test_headline_shape_step_bit_exact(env, ref, n_warm):
    """The step of BASELINE.json's headline config — HashGrid 16x2 T = 2^19, batch 2^18 (the view set stays small: the batch and the
    table size are what the kernels see) — against the oracle: same bit-exact bars as the small-shape test.  n_warm 2: the untrained
    grid (every cell occupied, ~16 batch sizes of samples before compaction, the inference cap max_samples is hit); n_warm 37: sparsified
    grid, automatic lane choice, prefetched rays."""
    batch = 1 << 18
    tb = _testbed(env, 0, True, HEADLINE_CFG)
    nr, n_written, nc = _check_step_against_oracle(tb, ref, env, n_warm, True, batch)
    assert n_written >= batch // 2 and nc > 0


@pytest.mark.parametrize("lanes", [1, 8, 32])
def test_sampler_with_cascades_and_cone_angle(env, ref, lanes):
    """aabb_scale 4: three occupancy cascades, cone angle 1/256 (exponential stepping through logf/expf) — the general,
    non-flat marcher path that fox (config 3) runs on.  Both sides evaluate logf/expf with the one explicitly rounded fp32
    implementation of include/ngp_b200_math.h, so the traversal is bit-exact here too: rays, per-ray counts, every coordinate."""
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(env["views"], aabb_scale=4)
    tb.reload_network_from_json(SMALL_CFG)
    rpb, max_inf, step = _step_and_capture(tb, ref, env, 3)
    aabb = ref.make_aabb((-1.5, -1.5, -1.5), (2.5, 2.5, 2.5))
    bf = tb.density_grid(1).cpu().numpy()
    o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, 1 + step * (1 << 32), env["images"], bf, 2, True, 1.0 / 256.0)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    assert int(c[0]) == int(o["ray_counter"]) and int(c[1]) == int(o["numsteps_counter"])
    nr = int(c[0])
    ns_gpu = tb.debug_tensor("numsteps_all", torch.int32, rpb).cpu().numpy().view(np.uint32)
    ref_idx = o["ray_indices"][:nr]; ref_cnt = o["numsteps"][:nr, 0]
    assert np.array_equal(ns_gpu[ref_idx], ref_cnt.astype(np.uint32))
    assert np.array_equal(tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32), ref_idx)
    n_written = int(ref_cnt.sum())
    gcoords = tb.debug_tensor("coords", torch.float32, n_written * 7).cpu().numpy().reshape(n_written, 7)
    assert np.array_equal(gcoords.view(np.uint32), o["coords"][:n_written].view(np.uint32))
    assert ns_gpu.max() > 0


@pytest.mark.parametrize("n_warm", [50, 62])
def test_walk_ahead_is_adopted_and_bit_exact(env, ref, n_warm):
    """walk_ahead (general marcher): the counting walk of step N+2 runs during step N+1 with a predicted ray count and only the emit phase
    follows once the control law has spoken.  A step whose rays came that way must equal the oracle bit for bit — same check as
    test_sampler_with_cascades_and_cone_angle, at step counts where neither that step nor the one before refreshes the occupancy grid —
    and the per-step traces must be those of a run with walk_ahead off."""
    pyngp = env["pyngp"]
    runs = []
    for wa in (True, False):
        tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
        if wa:
            tb.walk_ahead = True
        tb.set_nerf_dataset(env["views"], aabb_scale=4)
        tb.reload_network_from_json(SMALL_CFG)
        trace, hits = [], []
        for _ in range(n_warm):
            tb.train(BATCH)
            trace.append((tb.stats.rays_per_batch, tb.stats.measured_batch_size, tb.stats.measured_batch_size_before_compaction))
            hits.append(tb.walk_ahead_hits)
        rpb, max_inf, step = _step_and_capture(tb, ref, env, 0)
        runs.append(trace)
        if not wa:
            assert tb.walk_ahead_hits == 0 and tb.walk_ahead is False        # the default
            continue
        # the walk of step n_warm is adopted (emit phase launched) at the end of step n_warm - 1
        assert hits[-1] > hits[-2], (hits[-4:], tb.walk_ahead_misses)
        aabb = ref.make_aabb((-1.5, -1.5, -1.5), (2.5, 2.5, 2.5))
        bf = tb.density_grid(1).cpu().numpy()
        o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, 1 + step * (1 << 32), env["images"], bf, 2, True, 1.0 / 256.0)
        c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
        assert int(c[0]) == int(o["ray_counter"]) and int(c[1]) == int(o["numsteps_counter"])
        nr = int(c[0])
        ref_idx = o["ray_indices"][:nr]; ref_cnt = o["numsteps"][:nr, 0]
        ns_gpu = tb.debug_tensor("numsteps_all", torch.int32, rpb).cpu().numpy().view(np.uint32)
        assert np.array_equal(ns_gpu[ref_idx], ref_cnt.astype(np.uint32))
        assert np.array_equal(tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32), ref_idx)
        n_written = int(ref_cnt.sum())
        gcoords = tb.debug_tensor("coords", torch.float32, n_written * 7).cpu().numpy().reshape(n_written, 7)
        assert np.array_equal(gcoords.view(np.uint32), o["coords"][:n_written].view(np.uint32))
    assert runs[0] == runs[1]


def _perturbed_pose(xform, rng, angle, shift):
    """a pose `angle` radians and `shift` units away from xform (12 floats, column major)"""
    m = np.asarray(xform, np.float64).reshape(4, 3).T                       # 3 x 4
    w = rng.normal(size=3); w *= angle / np.linalg.norm(w)
    k = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    r = np.eye(3) + np.sin(angle) / angle * k + (1 - np.cos(angle)) / angle ** 2 * (k @ k)
    out = np.concatenate([r @ m[:, :3], (m[:, 3] + rng.normal(size=3) * shift)[:, None]], axis=1)
    return out.T.astype(np.float32).reshape(-1)


@pytest.mark.parametrize("moving", [False, True])
@pytest.mark.parametrize("lanes", [1, 8])
def test_rolling_shutter_rays_and_samples(env, ref, lanes, moving):
    """get_xform_given_rolling_shutter (common_device.cuh:401-408; the oracle's restatement is pinned bit for bit against the mount's own function,
    tests/test_oracle_pins.py): pixel time A + B u + C v + D motionblur_time picks a camera between the view's start and end pose for every ray.
    moving = False: start == end, non-zero shutter — glm's linear slerp branch, no transcendental, so rays and every sample are bit-exact.
    moving = True: poses 0.05 rad / 0.02 apart — acos / sin differ in the last ulp between libm and the device: rays to 2e-6."""
    rng = np.random.default_rng(11)
    views = []
    for i, v in enumerate(env["views"]):
        end = _perturbed_pose(v["xform"], rng, 0.05, 0.02) if moving else None
        views.append({**v, "xform_end": end, "rolling_shutter": (0.1, 0.2 * (i % 2), 0.5, 0.3)})
    images, keep = ref.images_from_views(views)
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(views, aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    rpb, max_inf, step = _step_and_capture(tb, ref, env, 2)
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    bf = tb.density_grid(1).cpu().numpy()
    o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, 1 + step * (1 << 32), images, bf, 0, True, 0.0)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    nr = int(o["ray_counter"])
    assert nr > 100
    if not moving:
        assert c[0] == o["ray_counter"] and c[1] == o["numsteps_counter"]
        assert np.array_equal(tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6), o["rays"][:nr])
        n_written = int(o["numsteps"][:nr, 0].sum())
        gcoords = tb.debug_tensor("coords", torch.float32, n_written * 7).cpu().numpy().reshape(n_written, 7)
        assert np.array_equal(gcoords, o["coords"][:n_written])
    else:
        assert abs(int(c[0]) - nr) <= 2                                   # a ray grazing the box may fall either side
        if int(c[0]) == nr:
            g = tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6)
            same = np.array_equal(tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32), o["ray_indices"][:nr])
            if same:
                np.testing.assert_allclose(g, o["rays"][:nr], rtol=0, atol=2e-6)
    # the shutter matters: the same step without it gives other rays
    plain, _ = ref.images_from_views(env["views"])
    o2 = ref.generate_training_samples(rpb, aabb, max_inf, 1337, 1 + step * (1 << 32), plain, bf, 0, True, 0.0)
    n2 = min(nr, int(o2["ray_counter"]))
    assert np.abs(o["rays"][:n2] - o2["rays"][:n2]).max() > (1e-3 if moving else 1e-8)


FOX_LENS = (0.0578421, -0.0805099, -0.000980296, 0.00015575)          # k1, k2, p1, p2 of data/nerf/fox/transforms.json:6-9
FOX_PRINCIPAL = (554.558 / 1080.0, 965.268 / 1920.0)                  # cx / w, cy / h (:10-13)


@pytest.mark.parametrize("lanes", [1, 8])
def test_opencv_lens_rays_and_samples_bit_exact(env, ref, lanes):
    """Config 3 (fox) trains every ray through iterative_opencv_lens_undistortion (common_device.cuh:289-377, up to 100 Newton
    iterations) with an off-centre principal point, 4 occupancy cascades (aabb_scale 8) and cone angle 1/256.  Same views as the other
    tests but declared with fox's lens: the ray origins / directions (undistorted on the GPU in sampler_classify_kernel), the cells
    mark_untrained_density_grid keeps (pos_to_uv applies the forward distortion), per-ray counts and every sample coordinate must equal
    the oracle's restatement bit for bit."""
    pyngp = env["pyngp"]
    views = [{**v, "lens_mode": 1, "lens_params": FOX_LENS, "principal_point": FOX_PRINCIPAL} for v in env["views"]]
    images, keep = ref.images_from_views(views)
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(views, aabb_scale=8)
    tb.reload_network_from_json(SMALL_CFG)
    assert tb.training_view(0)["lens_mode"] == 1
    rpb, max_inf, step = _step_and_capture(tb, ref, env, 2)
    NC = 128 ** 3
    grid0 = ref.mark_untrained(np.zeros(NC * 4, np.float32), images)
    grid = tb.density_grid(0).cpu().numpy()
    assert grid.size == NC * 4 and np.array_equal(grid == -1, grid0 == -1)
    aabb = ref.make_aabb((-3.5,) * 3, (4.5,) * 3)
    bf = tb.density_grid(1).cpu().numpy()
    o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, 1 + step * (1 << 32), images, bf, 3, True, 1.0 / 256.0)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    assert int(c[0]) == int(o["ray_counter"]) and int(c[1]) == int(o["numsteps_counter"])
    nr = int(c[0])
    assert nr > 0
    assert np.array_equal(tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32), o["ray_indices"][:nr])
    grays = tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6)
    assert np.array_equal(grays.view(np.uint32), o["rays"][:nr].view(np.uint32))
    # the lens did bend the rays: the same pixels through a pinhole give other directions
    o_pin = ref.generate_training_samples(rpb, aabb, max_inf, 1337, 1 + step * (1 << 32), env["images"], bf, 3, True, 1.0 / 256.0)
    assert not np.array_equal(o_pin["rays"][:8], o["rays"][:8])
    n_written = int(o["numsteps"][:nr, 0].sum())
    gcoords = tb.debug_tensor("coords", torch.float32, n_written * 7).cpu().numpy().reshape(n_written, 7)
    assert np.array_equal(gcoords.view(np.uint32), o["coords"][:n_written].view(np.uint32))


def test_marcher_variants_train_identically(env):
    """thread-per-ray, 8 and 32 lanes per ray: same samples and sample counts at every step of a 40-step run"""
    params = []
    for lanes, prefetch in ((1, False), (2, True), (4, True), (8, True), (32, True), (0, False), (0, True)):
        tb = _testbed(env, lanes, prefetch)
        for _ in range(40):
            tb.train(BATCH)
        params.append((tb.param_buffer(0).cpu().numpy().copy(), tb.stats.rays_per_batch, tb.stats.measured_batch_size))
    for p, r, m in params[1:]:
        assert (r, m) == params[0][1:]
        # samples are identical; hash-table gradients are summed with fp16 atomics whose order varies from run to run (for
        # any variant), which 40 Adam steps amplify on a few entries: compare the parameter vectors in aggregate
        assert np.linalg.norm(p - params[0][0]) <= 1e-2 * np.linalg.norm(params[0][0])


def test_device_control_law_trains_identically(env):
    """device_control_law (default on): rays_per_batch of the next step is ALSO computed on the device right behind the loss kernel and the
    next sampler is enqueued before the host has read the counters.  Same law, same integers: the per-step traces (rays_per_batch, compacted
    and un-compacted sample counts) must be identical with it off, the prefetch must actually be adopted, and a step captured while it is on
    must equal the oracle bit for bit (test_sampler_and_loss_bit_exact_compaction at n_warm 37 / 51 runs with it on)."""
    runs = []
    for dcl in (True, False):
        tb = _testbed(env)
        tb.device_control_law = dcl
        trace = []
        for _ in range(70):
            tb.train(BATCH)
            st = tb.stats
            trace.append((st.rays_per_batch, st.measured_batch_size, st.measured_batch_size_before_compaction))
        runs.append((trace, int(tb.prefetch_hits)))
    assert runs[0][0] == runs[1][0]
    assert runs[0][1] > 20 and runs[1][1] > 20, (runs[0][1], runs[1][1])


def test_deferred_step_finish_trains_identically(env):
    """deferred_step_finish (default on): when the next step's sampler already runs behind the device-side law, train() returns without
    waiting for the step's counters and the host half of the step's end (NerfCounters::update_after_training, testbed_nerf.cu:3583-3609) runs at
    the start of the next call.  Same arithmetic at a later moment: per-step traces, the running sums and the losses of every 16th step are
    identical with it off, with and without a view uploaded between the steps; training_step is current without forcing the wait."""
    runs = []
    for deferred in (True, False):
        tb = _testbed(env)
        tb.deferred_step_finish = deferred
        trace, losses = [], []
        for i in range(70):
            if i >= 30:
                tb.set_training_image(i % N_VIEWS, env["views"][i % N_VIEWS]["pixels"])      # the same pixels again: content unchanged
            tb.train(BATCH)
            assert tb.training_step == i + 1
            if i % 3 == 0 or i >= 60:                                                         # most steps are left pending on purpose
                st = tb.stats
                trace.append((i, st.rays_per_batch, st.measured_batch_size, st.measured_batch_size_before_compaction))
                losses.append(st.loss)
        totals = (tb.get_option("measured_batch_size_total"), tb.get_option("measured_batch_size_before_compaction_total"))
        runs.append((trace, totals, losses, int(tb.prefetch_hits)))
    assert runs[0][0] == runs[1][0] and runs[0][1] == runs[1][1]
    assert runs[0][3] > 20 and runs[0][3] == runs[1][3]
    np.testing.assert_allclose(runs[0][2], runs[1][2], rtol=2e-2)          # float atomics in the steps before: equal up to their order
    assert runs[0][1][0] > 60 * BATCH * 0.5 and all(np.isfinite(runs[0][2]))


def test_native_data_parallel_step_world_size_1(env):
    """the in-library data-parallel step (NCCL communicator bound at run time, counters + gradient all-reduce on the
    communication stream) with a single rank must train exactly like the plain step"""
    import ctypes as C
    _lib = importlib.import_module("instant-ngp_b200._lib")
    lib = _lib.lib()
    runs = []
    for native in (False, True):
        tb = _testbed(env)
        if native:
            ident = (C.c_uint8 * 128)()
            _lib.check(lib.ngp_comm_unique_id(ident, None))
            _lib.check(lib.ngp_testbed_comm_init(tb._h, ident, 0, 1, None))
        trace = []
        for _ in range(40):
            tb.train(BATCH)
            trace.append((tb.stats.rays_per_batch, tb.stats.measured_batch_size, tb.stats.measured_batch_size_before_compaction))
        runs.append((trace, tb.param_buffer(0).cpu().numpy().copy()))
    assert runs[0][0] == runs[1][0]
    assert np.linalg.norm(runs[0][1] - runs[1][1]) <= 1e-2 * np.linalg.norm(runs[0][1])


def test_set_training_image_replaces_the_targets(env):
    """nerf.training.set_image (python_api.cu:49-67): pixels uploaded after the dataset was set are what the loss sees — a testbed
    whose views were all replaced before the first step computes the same per-ray losses as one built from those pixels"""
    views2 = []
    for v in env["views"]:
        px = v["pixels"].clone()
        px[..., 0] = 255 - px[..., 0]                       # a different picture, same alpha
        views2.append({**v, "pixels": px})
    pyngp = env["pyngp"]
    a = pyngp.Testbed(pyngp.TestbedMode.Nerf); a.set_nerf_dataset(views2, aabb_scale=1); a.reload_network_from_json(SMALL_CFG)
    b = _testbed(env)
    for i, v in enumerate(views2):
        b.set_training_image(i, v["pixels"])
    c = _testbed(env)                                        # the original pictures: must differ
    out = []
    for tb in (a, b, c):
        tb.train(BATCH)
        out.append(tb.debug_tensor("loss", torch.float32, tb.stats.rays_per_batch if False else 4096).cpu().numpy().copy())
    np.testing.assert_allclose(out[1], out[0], rtol=1e-5, atol=1e-12)
    assert np.abs(out[2] - out[0]).max() > 1e-9
    with pytest.raises(Exception):
        b.set_training_image(len(views2), views2[0]["pixels"])


def test_rays_per_batch_control_law(env, ref):
    tb = _testbed(env)
    for _ in range(8):
        before = tb.stats
        tb.train(BATCH)
        after = tb.stats
        assert after.rays_per_batch == ref.lib().ref_next_rays_per_batch(before.rays_per_batch, BATCH, after.measured_batch_size)
    assert after.training_step == 8


def _psnr(tb, env, view):
    scene = env["scene"]
    v = env["views"][view]
    img = tb.render(RES, RES, spp=1, camera_matrix=v["xform"], focal_length=v["focal_length"])
    px = v["pixels"].numpy().astype(np.float32) / 255.0
    a = px[..., 3:4]
    lin = np.where(px[..., :3] <= 0.04045, px[..., :3] / 12.92, ((px[..., :3] + 0.055) / 1.055) ** 2.4) * a
    mse = float(((img[..., :3] - lin) ** 2).mean())
    return -10 * np.log10(max(mse, 1e-10)), img


def test_training_converges_and_renders(env):
    tb = _testbed(env)
    tb.snap_to_pixel_centers = True
    p0, _ = _psnr(tb, env, 0)
    losses = []
    for i in range(300):
        tb.train(BATCH)
        if i % 16 == 0:
            losses.append(tb.loss)
    p1, img = _psnr(tb, env, 0)
    assert losses[-1] < 0.5 * losses[0], losses
    assert p1 > p0 + 5 and p1 > 20, (p0, p1)
    assert np.isfinite(img).all() and img[..., 3].max() > 0.9
    # the run.py evaluation protocol (sRGB, black background, view's own camera): same scene, same ballpark as the linear PSNR above
    ev = tb.evaluate_training_views([0, 5], spp=2)
    assert ev["n"] == 2 and ev["min_psnr"] <= ev["psnr"] <= ev["max_psnr"] and ev["max_psnr"] > 20 and ev["psnr"] > 15, ev   # 300 small steps: floaters weigh more in sRGB
    assert tb.render_min_transmittance == pytest.approx(0.01) and tb.snap_to_pixel_centers     # knobs restored


def test_render_matches_oracle_and_tiles(env, ref):
    tb = _testbed(env)
    for _ in range(60):
        tb.train(BATCH)
    v = env["views"][3]
    W = H = 48
    f = env["scene"].focal_px(W)
    full = tb.render(W, H, spp=1, camera_matrix=v["xform"], focal_length=(f, f))
    top = tb.render(W, H, spp=1, camera_matrix=v["xform"], focal_length=(f, f), y0=0, y1=20)
    bot = tb.render(W, H, spp=1, camera_matrix=v["xform"], focal_length=(f, f), y0=20, y1=H)
    assert np.array_equal(np.concatenate([top, bot], 0), full)        # tile-sharded render == full render
    for world in (2, 3):                                              # interleaved rows (rank r: rows r, r + world, ...) == the same rows of the full render
        for r in range(world):
            part = tb.render(W, H, spp=1, camera_matrix=v["xform"], focal_length=(f, f), y0=r, row_stride=world)
            assert np.array_equal(part, full[r::world]), (world, r)
    rc = ref.RenderCfg()
    rc.res = (ref.C.c_int32 * 2)(W, H); rc.focal_length = (ref.f32 * 2)(f, f); rc.screen_center = (ref.f32 * 2)(0.5, 0.5)
    rc.camera = (ref.f32 * 12)(*[float(x) for x in v["xform"]])
    rc.render_aabb = ref.make_aabb((0, 0, 0), (1, 1, 1)); rc.train_aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    rc.cone_angle_constant = 0.0; rc.min_transmittance = 0.01; rc.max_mip = 0; rc.spp_index = 0; rc.snap_to_pixel_centers = 1
    rc.rgb_activation = 2; rc.density_activation = 3; rc.linear_colors = 0; rc.y0 = 0; rc.y1 = H
    params = tb.param_buffer(2).cpu().numpy()
    bf = tb.density_grid(1).cpu().numpy()
    rgba, depth = ref.render_nerf(rc, ref.nerf_cfg_from_json(SMALL_CFG), params, bf)
    assert np.abs(full - rgba).max() <= 2e-2, np.abs(full - rgba).max()
    assert np.abs(full - rgba).mean() <= 1e-3


def test_camera_setters_and_reset(env):
    """nerf.training.set_camera_intrinsics / set_camera_extrinsics (testbed_nerf.cu:2873-2919) and Testbed.reset"""
    tb, ref_tb = _testbed(env), _testbed(env)
    v0 = tb.training_view(2)
    x = np.asarray(v0["xform"], np.float64).reshape(4, 3).T            # NGP 3x4
    nerf = x[[2, 0, 1], :].copy(); nerf[:, 1] *= -1; nerf[:, 2] *= -1; nerf[:, 3] = (nerf[:, 3] - 0.5) / 0.33      # ngp_matrix_to_nerf
    tb.set_camera_extrinsics(2, np.eye(4)[:3], convert_to_ngp=False)
    assert np.array_equal(tb.training_view(2)["xform"], np.eye(4, dtype=np.float32)[:3, :4].T.reshape(-1))
    tb.set_camera_extrinsics(2, nerf, convert_to_ngp=True)
    np.testing.assert_allclose(tb.training_view(2)["xform"], v0["xform"], atol=1e-6)
    tb.set_camera_extrinsics(2, x, convert_to_ngp=False)                # exact restore
    assert np.array_equal(tb.training_view(2)["xform"], v0["xform"])
    tb.set_camera_intrinsics(2, fx=100.0, cx=24.0, cy=-0.25, k1=0.1, p2=0.01)
    v = tb.training_view(2)
    assert v["focal_length"] == (100.0, 100.0) and v["principal_point"] == (24.0 / RES, 0.25) and v["lens_mode"] == 1
    assert v["lens_params"] == (np.float32(0.1), 0.0, 0.0, np.float32(0.01))
    tb.set_camera_intrinsics(2, fx=v0["focal_length"][0], fy=v0["focal_length"][1], cx=-0.5, cy=-0.5)
    v = tb.training_view(2)
    assert v["lens_mode"] == 0 and v["focal_length"] == v0["focal_length"] and v["principal_point"] == (0.5, 0.5)
    tb.set_camera_intrinsics(N_VIEWS + 5, fx=1.0)                       # out of range: ignored
    # the device copy is what the sampler reads: with everything restored the first step is identical to an untouched testbed's
    tb.train(BATCH); ref_tb.train(BATCH)
    n = ref_tb.stats.rays_per_batch
    assert torch.equal(tb.debug_tensor("rays", torch.float32, n * 6), ref_tb.debug_tensor("rays", torch.float32, n * 6))
    # (parameters: identical up to the entries whose fp16-atomic gradient sum cancels to exactly 0 in one summation order only, see below)
    assert (tb.param_buffer(0) != ref_tb.param_buffer(0)).float().mean().item() <= 1e-5
    # reset: step counter, counters and parameters as after reload_network
    fresh = _testbed(env)
    for _ in range(5):
        tb.train(BATCH)
    tb.reset()
    assert tb.training_step == 0 and tb.stats.rays_per_batch == 4096
    assert torch.equal(tb.param_buffer(0), fresh.param_buffer(0)) and float(tb.density_grid(0).abs().sum()) == 0.0
    tb.train(BATCH); fresh.train(BATCH)
    # one step from the same state: the same samples, but the hash-table gradients are summed with fp16 atomics in an order that changes from
    # run to run — an entry whose contributions cancel to exactly 0 in one order is skipped by the sparse Adam rule and stepped by +-lr in
    # another (seen once in ~10 runs of this file).  Everything else must be identical.
    differ = (tb.param_buffer(0) != fresh.param_buffer(0)).float().mean().item()
    assert differ <= 1e-5, differ
