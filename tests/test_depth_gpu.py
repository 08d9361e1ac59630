"""Depth supervision (nerf.training.depth_supervision_lambda / depth_loss_type, python_api.cu:616,635): views may carry a depth
map (NerfDataset::set_training_image's depth half, nerf_loader.cu:943-959; "integer_depth_scale" + "depth_path" in transforms.json, :471-473,609-621);
the loss kernel composites the depth of every ray, compares it with the map at the ray's pixel and adds lambda * dL_depth/d depth to the density
gradient (testbed_nerf.cu:1743-1752,1848-1856,1926-1939).  Against the oracle on the GPU's own network output, in every marcher / loss-kernel variant."""
import importlib
import json

import numpy as np
import pytest
import torch

from .test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG, _step_and_capture
from .test_nerf_loader_gpu import _ngp_to_nerf

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(ngp, ref):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    rng = np.random.default_rng(17)
    yy, xx = np.mgrid[0:RES, 0:RES].astype(np.float32) / RES
    depth = []                                             # what the library should hold per view (scene units), None = no map
    raw = []                                               # (host array, depth_scale) as handed over
    for i in range(N_VIEWS):
        if i % 3 == 1:                                     # (view 1: early in training only the rays of views 0 and 1 fit into the batch)
            depth.append(None); raw.append(None)
        elif i % 3 == 0:                                   # float map, scale 0.5; a tenth of the pixels carry no depth (0)
            d = (1.2 + 0.8 * np.sin(6.0 * xx + i) * np.cos(5.0 * yy)).astype(np.float32)
            d[rng.random((RES, RES)) < 0.1] = 0.0
            raw.append((d, 0.5)); depth.append((d * np.float32(0.5)).astype(np.float32))
        else:                                              # 16-bit map with an integer scale, as a depth PNG gives it
            d = (rng.integers(200, 1200, (RES, RES))).astype(np.uint16)
            raw.append((d, 1.0 / 1000.0)); depth.append((d.astype(np.float32) * np.float32(1.0 / 1000.0)).astype(np.float32))
    images, keep = ref.images_from_views(views)
    for i, d in enumerate(depth):
        if d is not None:
            images[i].depth = ref.P(d).value
    return dict(pyngp=pyngp, scene=scene, views=views, images=images, keep=(keep, depth), depth=depth, raw=raw)


def _testbed(env, lanes):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    for i, r in enumerate(env["raw"]):
        if r is not None:
            tb.set_training_depth(i, r[0], r[1])
    return tb


def _oracle(ref, env, tb, rpb, max_inf, step, lam, loss_type):
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    bf = tb.density_grid(1).cpu().numpy()
    skip = 1 + step * (1 << 32)
    o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
    nr = int(o["ray_counter"]); n_written = int(o["numsteps"][:nr, 0].sum())
    mlp_out = tb.debug_tensor("mlp_out", torch.float16, n_written * 4).cpu().numpy().reshape(n_written, 4)
    mean_density = float(tb.debug_tensor("mean_density", torch.float32, 1).cpu().numpy()[0])
    cfg = ref.LossCfg(4, 2, 3, 1, 0, 1, (ref.f32 * 3)(0, 0, 0), 128.0, 0.1, mean_density)
    cfg.depth_supervision_lambda = lam; cfg.depth_loss_type = loss_type
    lo = ref.compute_loss(rpb, aabb, 1337, skip, BATCH, nr, cfg, env["images"], mlp_out, o["ray_indices"], o["rays"], o["numsteps"], o["coords"])
    return lo, nr


@pytest.mark.parametrize("lanes,n_warm,loss_type", [(1, 4, 1), (8, 4, 1), (32, 4, 1), (0, 37, 1), (8, 20, 0), (8, 20, 4)])
def test_depth_term_of_the_loss_gradient(env, ref, lanes, n_warm, loss_type):
    tb = _testbed(env, lanes)
    pyngp = env["pyngp"]
    assert tb.nerf.training.depth_supervision_lambda == 0.0 and tb.nerf.training.depth_loss_type == pyngp.LossType.L1
    lam = 0.7
    tb.nerf.training.depth_supervision_lambda = lam
    tb.nerf.training.depth_loss_type = pyngp.LossType(loss_type)
    rpb, max_inf, step = _step_and_capture(tb, ref, env, n_warm)
    lo, nr = _oracle(ref, env, tb, rpb, max_inf, step, lam, loss_type)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    assert c[2] == lo["compacted_counter"]
    nc = int(min(c[2], BATCH))
    gdl = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    scale = np.abs(rdl).max()
    err = np.abs(gdl[:nc] - rdl).max()
    assert err <= 2e-3 * scale + 1e-7, (err, scale)
    # the term is there: the same step without depth maps has other density gradients, the colour gradients are untouched
    plain, _ = _oracle(ref, env, tb, rpb, max_inf, step, 0.0, loss_type)
    pdl = plain["dloss"][:nc].astype(np.float32)
    assert np.array_equal(pdl[:, :3], rdl[:, :3])
    assert np.abs(pdl[:, 3] - rdl[:, 3]).max() > 50 * max(err, 1e-7)
    # rays of the views without a map (every third; several hundred samples per ray at this stage: the batch holds the rays of views 0 and 1) get no term
    ns = lo["numsteps"][:nr]
    img_of_ray = (np.arange(rpb, dtype=np.int64) * N_VIEWS // rpb)          # image_idx of ray slot i (testbed_nerf.cu:1317-1326)
    ray_idx = tb.debug_tensor("ray_indices", torch.int32, nr).cpu().numpy().view(np.uint32)
    no_map = (img_of_ray[ray_idx] % 3 == 1)
    rows = np.concatenate([np.arange(b, min(b + n, nc)) for (n, b), f in zip(ns, no_map) if f and n > 0 and b < nc] or [np.zeros(0, np.int64)]).astype(np.int64)
    assert rows.size > 0, (int(no_map.sum()), nr, nc, rpb, ns[no_map][:5].tolist(), ray_idx[:5].tolist(), ray_idx[-5:].tolist())
    assert np.array_equal(pdl[rows], rdl[rows])


def test_depth_maps_can_be_replaced_and_dropped(env, ref):
    tb = _testbed(env, 8)
    tb.nerf.training.depth_supervision_lambda = 0.7
    for i, r in enumerate(env["raw"]):
        if r is not None:
            tb.set_training_depth(i, None, -1.0)           # depth_scale < 0: the map is dropped (nerf_loader.cu:957-959)
    rpb, max_inf, step = _step_and_capture(tb, ref, env, 6)
    lo, nr = _oracle(ref, env, tb, rpb, max_inf, step, 0.0, 1)
    nc = int(min(lo["compacted_counter"], BATCH))
    gdl = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    assert np.abs(gdl[:nc] - rdl).max() <= 2e-3 * np.abs(rdl).max() + 1e-7
    with pytest.raises(RuntimeError, match="Invalid frame index"):
        tb.set_training_depth(N_VIEWS, np.zeros((RES, RES), np.float32), 1.0)


def test_depth_maps_from_transforms_json(env, ref, tmp_path):
    """"integer_depth_scale" + per-frame "depth_path" (16-bit PNG): loaded maps = pixel * integer_depth_scale * scene scale"""
    import cv2
    pyngp, scene = env["pyngp"], env["scene"]
    views = env["views"][:4]
    (tmp_path / "images").mkdir(); (tmp_path / "depth").mkdir()
    frames, maps = [], []
    rng = np.random.default_rng(3)
    for i, v in enumerate(views):
        cv2.imwrite(str(tmp_path / "images" / f"{i:03d}.png"), np.ascontiguousarray(v["pixels"].numpy()[..., [2, 1, 0, 3]]))
        fr = {"file_path": f"images/{i:03d}.png", "transform_matrix": _ngp_to_nerf(v["xform"])}
        if i != 2:                                         # one frame without a depth file: a map of zeros
            d16 = rng.integers(1000, 60000, (RES, RES)).astype(np.uint16)
            cv2.imwrite(str(tmp_path / "depth" / f"{i:03d}.png"), d16)
            fr["depth_path"] = f"depth/{i:03d}.png"
            maps.append(d16)
        else:
            maps.append(None)
        frames.append(fr)
    doc = {"camera_angle_x": scene.CAMERA_ANGLE_X, "aabb_scale": 1, "integer_depth_scale": 1.0 / 20000.0, "frames": frames}
    (tmp_path / "transforms.json").write_text(json.dumps(doc))
    tb = pyngp.Testbed()
    tb.load_training_data(str(tmp_path / "transforms.json"))
    tb.reload_network_from_json(SMALL_CFG)
    tb.nerf.training.depth_supervision_lambda = 1.0
    scale = np.float32(np.float32(1.0 / 20000.0) * np.float32(0.33))                     # NERF_SCALE, the loader's default scene scale
    depth = [np.zeros((RES, RES), np.float32) if m is None else (m.astype(np.float32) * scale).astype(np.float32) for m in maps]
    # the oracle works on the views exactly as the loader produced them (poses go through the NeRF <-> NGP conversion: equal to ~1e-6, not bitwise)
    loaded = [{**v, **{k: tb.training_view(i)[k] for k in ("xform", "focal_length", "principal_point")}} for i, v in enumerate(views)]
    images, keep = ref.images_from_views(loaded)
    for i, d in enumerate(depth):
        images[i].depth = ref.P(d).value
    e2 = dict(env, views=loaded, images=images, keep=(keep, depth))
    rpb, max_inf, step = _step_and_capture(tb, ref, e2, 5)
    lo, nr = _oracle(ref, e2, tb, rpb, max_inf, step, 1.0, 1)
    nc = int(min(lo["compacted_counter"], BATCH))
    gdl = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    assert np.abs(gdl[:nc] - rdl).max() <= 2e-3 * np.abs(rdl).max() + 1e-7
    # without "integer_depth_scale" the depth files are not even looked at
    doc.pop("integer_depth_scale")
    (tmp_path / "transforms.json").write_text(json.dumps(doc))
    tb2 = pyngp.Testbed()
    tb2.load_training_data(str(tmp_path / "transforms.json"))
    tb2.reload_network_from_json(SMALL_CFG)
    tb2.nerf.training.depth_supervision_lambda = 1.0
    rpb, max_inf, step = _step_and_capture(tb2, ref, e2, 5)
    lo, nr = _oracle(ref, e2, tb2, rpb, max_inf, step, 0.0, 1)
    nc = int(min(lo["compacted_counter"], BATCH))
    gdl = tb2.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    assert np.abs(gdl[:nc] - rdl).max() <= 2e-3 * np.abs(rdl).max() + 1e-7
