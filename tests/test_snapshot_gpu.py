"""Snapshots (Testbed.save_snapshot / load_snapshot, testbed.cu:4873-5057): the file is read back with an independent
MessagePack / zlib implementation (the `msgpack` package) and compared with the device buffers; a fresh testbed that loads it
renders bit-identically and trains on; optimizer state round-trips exactly."""
import importlib
import zlib

import msgpack
import numpy as np
import pytest
import torch

from tests.test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(ngp):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    return dict(pyngp=pyngp, views=scene.make_views(N_VIEWS, RES, RES))


def _trained(env, steps=40):
    tb = env["pyngp"].Testbed(env["pyngp"].TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    for _ in range(steps):
        tb.train(BATCH)
    return tb


def _render(tb, env):
    v = env["views"][3]
    return tb.render(RES, RES, spp=1, camera_matrix=v["xform"], focal_length=v["focal_length"])


@pytest.mark.parametrize("name, compress", [("a.ingp", True), ("b.ingp", False), ("c.msgpack", True)])
def test_snapshot_file_format_and_reload(env, tmp_path, name, compress):
    tb = _trained(env)
    path = tmp_path / name
    tb.save_snapshot(path, include_optimizer_state=False, compress=compress)
    raw = path.read_bytes()
    if name.endswith(".ingp"):
        raw = zlib.decompress(raw)                      # a plain zlib stream (zstr::ostream)
    cfg = msgpack.unpackb(raw, raw=False)
    # the network config is embedded unchanged, the snapshot object carries what testbed.cu:4873-4936 writes
    for k in ("loss", "optimizer", "encoding", "network", "rgb_network"):
        assert cfg[k] == SMALL_CFG[k]
    snap = cfg["snapshot"]
    assert snap["version"] == 1 and snap["mode"] == "nerf" and snap["density_grid_size"] == 128
    assert snap["n_params"] == tb.n_params() and snap["params_type"] == "__half"
    st = tb.stats
    assert snap["training_step"] == st.training_step == 40 and abs(snap["loss"] - st.loss) < 1e-7
    assert snap["nerf"]["rgb"] == {"rays_per_batch": st.rays_per_batch, "measured_batch_size": st.measured_batch_size,
                                   "measured_batch_size_before_compaction": st.measured_batch_size_before_compaction}
    assert snap["nerf"]["aabb_scale"] == 1 and snap["nerf"]["dataset"]["n_images"] == N_VIEWS
    x0 = np.asarray(snap["nerf"]["dataset"]["xforms"][0]["start"], np.float32)          # 3 rows of 4 (json_binding.h:45-66)
    assert np.array_equal(x0.T.reshape(-1), np.asarray(env["views"][0]["xform"], np.float32))
    params = np.frombuffer(snap["params_binary"], np.float16)
    assert np.array_equal(params, tb.param_buffer(2).cpu().numpy())                      # the (EMA) inference parameters
    grid = np.frombuffer(snap["density_grid_binary"], np.float16)
    assert np.array_equal(grid, tb.density_grid(0).half().cpu().numpy())
    assert "optimizer" not in snap

    # a fresh testbed with no data set and no network: everything needed to render comes from the file
    tb2 = env["pyngp"].Testbed(env["pyngp"].TestbedMode.Nerf)
    tb2.load_file(path)
    assert tb2.training_step == 40 and tb2.n_params() == tb.n_params()
    assert torch.equal(tb2.param_buffer(2), tb.param_buffer(2)) and torch.equal(tb2.param_buffer(1), tb.param_buffer(2))
    assert torch.equal(tb2.param_buffer(0), tb.param_buffer(2).float())                  # Trainer::set_params of fp16 weights
    assert torch.equal(tb2.density_grid(1), tb.density_grid(1))                          # bitfield rebuilt from the fp16 grid
    assert np.array_equal(_render(tb2, env), _render(tb, env))
    with pytest.raises(RuntimeError, match="training data"):
        tb2.train(BATCH)
    # with the data set present, training carries on from the snapshot
    tb3 = env["pyngp"].Testbed(env["pyngp"].TestbedMode.Nerf)
    tb3.set_nerf_dataset(env["views"], aabb_scale=1)
    tb3.load_snapshot(path)
    assert tb3.stats.rays_per_batch == st.rays_per_batch
    for _ in range(17):
        tb3.train(BATCH)
    assert tb3.training_step == 57 and np.isfinite(tb3.loss) and tb3.loss < 2 * st.loss


def test_snapshot_with_optimizer_state_is_exact(env, tmp_path):
    tb = _trained(env, 24)
    path = tmp_path / "opt.ingp"
    tb.save_snapshot(path, include_optimizer_state=True)
    snap = msgpack.unpackb(zlib.decompress(path.read_bytes()), raw=False)["snapshot"]
    adam = snap["optimizer"]["nested"]["nested"]                                          # Ema{ExponentialDecay{Adam}}
    assert adam["current_step"] == 24 and len(adam["first_moments_binary"]) == 4 * tb.n_params()
    steps = np.frombuffer(adam["param_steps_binary"], np.uint32)
    assert steps.max() == 24 and steps.min() == 0                                         # untouched hash entries never stepped
    tb2 = env["pyngp"].Testbed(env["pyngp"].TestbedMode.Nerf)
    tb2.set_nerf_dataset(env["views"], aabb_scale=1)
    tb2.load_snapshot(path)
    for which in (0, 1, 2):
        assert torch.equal(tb2.param_buffer(which), tb.param_buffer(which))
    path2 = tmp_path / "opt2.ingp"
    tb2.save_snapshot(path2, include_optimizer_state=True)
    assert zlib.decompress(path2.read_bytes()) == zlib.decompress(path.read_bytes())     # save(load(x)) == x, byte for byte
    assert abs(tb2.stats.learning_rate - tb.stats.learning_rate) < 1e-12


def test_snapshot_errors(env, tmp_path):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    with pytest.raises(RuntimeError, match="does not exist"):
        tb.load_snapshot(tmp_path / "missing.ingp")
    with pytest.raises(RuntimeError, match="no network"):
        tb.save_snapshot(tmp_path / "x.ingp")
    (tmp_path / "cfg.msgpack").write_bytes(msgpack.packb(SMALL_CFG))
    with pytest.raises(RuntimeError, match="does not contain a snapshot"):
        tb.load_snapshot(tmp_path / "cfg.msgpack")
    (tmp_path / "old.msgpack").write_bytes(msgpack.packb({**SMALL_CFG, "snapshot": {"version": 0}}))
    with pytest.raises(RuntimeError, match="old format"):
        tb.load_snapshot(tmp_path / "old.msgpack")
    (tmp_path / "bad.ingp").write_bytes(b"not a zlib stream")
    with pytest.raises(RuntimeError, match="zlib"):
        tb.load_snapshot(tmp_path / "bad.ingp")
    img = pyngp.Testbed(pyngp.TestbedMode.Image)
    t = _trained(env, 2)
    t.save_snapshot(tmp_path / "n.msgpack")
    # a C-ABI testbed has one mode for life and refuses a snapshot of another; the binding does what Testbed::load_snapshot does
    # (testbed.cu:4950-4955): it switches to the snapshot's mode
    _lib = importlib.import_module("instant-ngp_b200._lib")
    assert _lib.lib().ngp_testbed_load_snapshot(img._h, None, str(tmp_path / "n.msgpack").encode()) != 0
    assert "mode" in _lib.lib().ngp_last_error().decode()
    img.load_snapshot(tmp_path / "n.msgpack")
    assert img.mode == pyngp.TestbedMode.Nerf and img.training_step == t.training_step


def test_snapshot_carries_view_state_and_trained_camera_offsets(env, tmp_path):
    """testbed.cu:4893-4894,4902-4913 / :5006-5031,5043-5054: the testbed's own camera, exposure and up vector, and the per-view pose offsets of
    optimize_extrinsics, which are stored beside the poses as given and applied again on load"""
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.nerf.training.optimize_extrinsics = True
    tb.nerf.training.n_steps_between_cam_updates = 4
    for _ in range(40):
        tb.train(BATCH)
    tb.exposure = 0.75
    tb.up_dir = [0.0, 0.0, 1.0]
    tb.set_camera_to_training_view(5)
    tb.zoom = 1.25; tb.screen_center = [0.4, 0.55]; tb.fov_axis = 0; tb.fov = 37.0
    poses = [tb.get_camera_extrinsics(i, convert_to_ngp=False) for i in range(N_VIEWS)]
    moved = max(np.abs(p - np.asarray(v["xform"], np.float32).reshape(4, 3).T).max() for p, v in zip(poses, env["views"]))
    assert moved > 1e-4
    path = tmp_path / "cams.msgpack"
    tb.save_snapshot(path)
    snap = msgpack.unpackb(path.read_bytes(), raw=False)["snapshot"]
    # the file: poses as given, offsets beside them, view state
    for i, v in enumerate(env["views"]):
        given = np.asarray(snap["nerf"]["dataset"]["xforms"][i]["start"], np.float32).T.reshape(-1)
        assert np.array_equal(given, np.asarray(v["xform"], np.float32))
    assert len(snap["nerf"]["cam_pos_offset"]) == N_VIEWS and snap["nerf"]["cam_rot_offset"][0]["iter"] == 10
    assert max(abs(x) for o in snap["nerf"]["cam_rot_offset"] for x in o["variable"]) > 1e-5
    assert abs(snap["exposure"] - 0.75) < 1e-7 and snap["up_dir"] == [0.0, 0.0, 1.0]
    assert snap["camera"]["fov_axis"] == 0 and abs(snap["camera"]["zoom"] - 1.25) < 1e-7 and len(snap["camera"]["matrix"]) in (3, 4)

    def check(other):
        assert abs(other.exposure - 0.75) < 1e-7 and list(other.up_dir) == [0.0, 0.0, 1.0]
        np.testing.assert_array_equal(other.camera_matrix, tb.camera_matrix)
        assert other.fov_axis == 0 and abs(other.zoom - 1.25) < 1e-7 and abs(other.fov - 37.0) < 1e-4
        np.testing.assert_allclose(other.screen_center, [0.4, 0.55], atol=1e-7)
        np.testing.assert_array_equal(other.relative_focal_length, tb.relative_focal_length)
        for i in range(N_VIEWS):
            np.testing.assert_array_equal(other.get_camera_extrinsics(i, convert_to_ngp=False), poses[i])

    fresh = pyngp.Testbed()                                  # no data set: views, offsets and camera all come from the file
    fresh.load_snapshot(path)
    check(fresh)
    same = pyngp.Testbed(pyngp.TestbedMode.Nerf)             # the same data set already loaded: the offsets apply to it
    same.set_nerf_dataset(env["views"], aabb_scale=1)
    same.load_snapshot(path)
    check(same)
    same.nerf.training.optimize_extrinsics = True
    for _ in range(8):
        same.train(BATCH)                                    # and training goes on from the trained poses
    assert np.isfinite(same.loss)
    other_data = pyngp.Testbed(pyngp.TestbedMode.Nerf)       # another data set: its own poses stay as they are
    shifted = [{**v, "xform": np.asarray(v["xform"], np.float32) + np.float32([0] * 9 + [0.01, 0, 0])} for v in env["views"]]
    other_data.set_nerf_dataset(shifted, aabb_scale=1)
    other_data.load_snapshot(path)
    for i, v in enumerate(shifted):
        np.testing.assert_array_equal(other_data.get_camera_extrinsics(i, convert_to_ngp=False), np.asarray(v["xform"], np.float32).reshape(4, 3).T)
