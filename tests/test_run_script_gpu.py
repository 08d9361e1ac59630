"""The flow of scripts/run.py against the binding (run.py:88-206,219-245): a Testbed without a mode, root_dir with configs/<mode>/ holding
commented configs and "parent" children, load_file on a scene directory, the frame() loop, save_snapshot, load_snapshot into an empty
testbed, reloading the test transforms into the trained testbed; the default network when none was given; Image mode from a file."""
import importlib
import json

import numpy as np
import pytest
import torch

from tests.test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG
from tests.test_nerf_loader_gpu import _ngp_to_nerf

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pyngp(ngp):
    return importlib.import_module("instant-ngp_b200.pyngp")


@pytest.fixture(scope="module")
def views(ngp):
    scene = importlib.import_module("instant-ngp_b200.scene")
    return scene.make_views(N_VIEWS, RES, RES)


def _write_scene(root, views, name="transforms.json"):
    import cv2
    scene = importlib.import_module("instant-ngp_b200.scene")
    (root / "images").mkdir(parents=True, exist_ok=True)
    frames = []
    for i, v in enumerate(views):
        cv2.imwrite(str(root / "images" / f"{i:04d}.png"), np.ascontiguousarray(v["pixels"].numpy()[..., [2, 1, 0, 3]]))
        frames.append({"file_path": f"images/{i:04d}.png", "transform_matrix": _ngp_to_nerf(v["xform"])})
    (root / name).write_text(json.dumps({"camera_angle_x": scene.CAMERA_ANGLE_X, "aabb_scale": 1, "frames": frames}))
    return root


def _write_configs(root):
    """configs/nerf/base.json in the reference's style (comments, tabs) and a child that patches one key of it"""
    d = root / "configs" / "nerf"
    d.mkdir(parents=True)
    base = json.dumps(SMALL_CFG if isinstance(SMALL_CFG, dict) else json.loads(SMALL_CFG), indent="\t")
    base = base.replace('"encoding": {', '// the position encoding\n\t"encoding": {\n\t\t/* "otype": "Frequency", */', 1)
    (d / "base.json").write_text(base)
    (d / "tiny.json").write_text('{\n\t"parent": "base.json",\n\t"encoding": {\n\t\t"log2_hashmap_size": 12 // a patch of one key\n\t}\n}\n')
    (d / "tinier.json").write_text('{"parent": "tiny.json", "network": {"n_neurons": 64}, "loss": {"otype": "L2"}}')
    return root


def test_run_py_main_flow(pyngp, views, tmp_path):
    root = _write_configs(tmp_path / "root")
    scene_dir = _write_scene(tmp_path / "scene", views)
    cfg = SMALL_CFG if isinstance(SMALL_CFG, dict) else json.loads(SMALL_CFG)

    testbed = pyngp.Testbed()
    testbed.root_dir = str(root)
    testbed.load_file(str(scene_dir))                               # run.py:93-95: a directory is training data, and the mode follows it
    assert testbed.mode == pyngp.TestbedMode.Nerf and testbed.nerf.training.dataset.n_images == N_VIEWS
    ds = testbed.nerf.training.dataset                                # NerfDataset members (python_api.cu:594-610)
    assert abs(ds.scale - 0.33) < 1e-7 and np.allclose(ds.offset, 0.5) and ds.aabb_scale == 1 and len(ds.paths) == N_VIEWS
    assert np.allclose(ds.transforms[3].start, np.asarray(views[3]["xform"], np.float32).reshape(4, 3).T, atol=2e-6)
    assert np.array_equal(ds.transforms[3].start, ds.transforms[3].end) and ds.metadata[3].rolling_shutter == [0.0, 0.0, 0.0, 0.0]
    testbed.reload_network_from_file("tinier.json")                 # resolved under <root_dir>/configs/nerf/, two parents up
    merged = json.loads(json.dumps(cfg)); merged["encoding"]["log2_hashmap_size"] = 12; merged["loss"] = {"otype": "L2"}
    explicit = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    explicit.reload_network_from_json(merged)
    assert testbed.n_params() == explicit.n_params()
    plain = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    plain.reload_network_from_json(cfg)
    assert testbed.n_params() < plain.n_params()                    # the child's table size is in force
    assert testbed.nerf.training.loss_type == pyngp.LossType.L2     # and its loss
    testbed.reload_network_from_file("base.json")
    assert testbed.n_params() == plain.n_params()
    testbed.reload_network_from_json({"parent": "tiny.json", "loss": {"otype": "L1"}}, str(root / "configs" / "nerf") + "/")
    assert testbed.n_params() == explicit.n_params() and testbed.nerf.training.loss_type == pyngp.LossType.L1
    with pytest.raises(RuntimeError, match="does not exist"):
        testbed.reload_network_from_file("no_such_config.json")
    testbed.reload_network_from_file(str(root / "configs" / "nerf" / "tiny.json"))      # a path that exists is taken as it is

    # run.py:131-170
    testbed.nerf.sharpen = 0.0
    testbed.exposure = 0.0
    testbed.shall_train = True
    testbed.nerf.render_with_lens_distortion = True
    testbed.nerf.training.near_distance = 0.05
    testbed.training_batch_size = BATCH
    n_steps, old_training_step = 300, 0
    while testbed.frame():                                          # run.py:183-202
        if testbed.want_repl():
            raise AssertionError("no REPL without a window")
        if testbed.training_step >= n_steps:
            break
        assert testbed.training_step > old_training_step
        old_training_step = testbed.training_step
    assert testbed.training_step == n_steps and np.isfinite(testbed.loss)
    snap = tmp_path / "out.ingp"
    testbed.save_snapshot(str(snap), False)

    # run.py:219-245 on the trained testbed: the test transforms are loaded into it, the network stays
    testbed.background_color = [0.0, 0.0, 0.0, 1.0]
    testbed.snap_to_pixel_centers = True
    testbed.nerf.render_min_transmittance = 1e-4
    testbed.shall_train = False
    testbed.set_camera_to_training_view(0)
    before = testbed.render(RES, RES, 1, True)
    _write_scene(tmp_path / "scene", views[:3], "transforms_test.json")
    testbed.load_training_data(str(scene_dir / "transforms_test.json"))
    assert testbed.nerf.training.dataset.n_images == 3 and testbed.training_step == n_steps
    res = testbed.nerf.training.dataset.metadata[0].resolution
    testbed.render_ground_truth = True
    testbed.set_camera_to_training_view(0)
    ref_image = testbed.render(res[0], res[1], 1, True)
    testbed.render_ground_truth = False
    image = testbed.render(res[0], res[1], 1, True)
    np.testing.assert_array_equal(image, before)
    mse = float(np.mean((image[..., :3] - ref_image[..., :3]) ** 2))
    assert mse < 0.05, mse

    # a second, empty testbed takes its mode from the snapshot (testbed.cu:4950-4955)
    other = pyngp.Testbed()
    other.background_color = [0.0, 0.0, 0.0, 1.0]                   # set before the mode exists: kept
    other.snap_to_pixel_centers = True
    other.load_file(str(snap))
    assert other.mode == pyngp.TestbedMode.Nerf and other.training_step == n_steps
    assert list(other.background_color) == [0.0, 0.0, 0.0, 1.0] and other.snap_to_pixel_centers is True
    other.nerf.render_min_transmittance = 1e-4
    other.set_camera_to_training_view(0)
    np.testing.assert_allclose(other.render(RES, RES, 1, True), before, atol=2e-3)


def test_default_network_when_none_was_given(pyngp, views):
    """testbed.cu:4297-4306: training data without a network -> "base.json"; with no configs directory the built-in hyper-parameters"""
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.root_dir = "/nonexistent_root"
    tb.set_nerf_dataset(views, aabb_scale=1)
    for _ in range(20):
        tb.train(BATCH)
    explicit = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    explicit.reload_network_from_json(pyngp.builtin_base_config(pyngp.TestbedMode.Nerf))
    assert tb.n_params() == explicit.n_params() and tb.nerf.training.loss_type == pyngp.LossType.Huber
    assert tb.training_step == 20 and np.isfinite(tb.loss)
    tb.reload_network_from_file()                                   # "": the remembered one again = a reset
    assert tb.training_step == 0


def _srgb_to_linear(s):
    return np.where(s <= 0.04045, s / 12.92, ((s + 0.055) / 1.055) ** 2.4)


IMAGE_CFG = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15},
             "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16},
             "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}}


def test_image_mode_from_a_file(pyngp, views, tmp_path):
    import cv2
    rng = np.random.default_rng(5)
    h, w = 48, 64
    yy, xx = np.mgrid[0:h, 0:w]
    rgba = np.stack([(xx * 4) % 256, (yy * 5) % 256, ((xx + yy) * 3) % 256, 255 - (xx * 2) % 200], axis=2).astype(np.uint8)
    rgba[rng.random((h, w)) < 0.1] = (255, 255, 255, 0)
    cv2.imwrite(str(tmp_path / "picture.png"), np.ascontiguousarray(rgba[..., [2, 1, 0, 3]]))

    tb = pyngp.Testbed()
    tb.exposure = 0.5
    tb.load_training_data(str(tmp_path / "picture.png"))           # anything that is not a scene, mesh or volume is an image to learn
    assert tb.mode == pyngp.TestbedMode.Image and tb.exposure == 0.5
    got = tb.debug_tensor("image", torch.float32, h * w * 4).cpu().numpy().reshape(h, w, 4)
    alpha = rgba[..., 3:4].astype(np.float32) / 255.0
    want = np.concatenate([_srgb_to_linear(rgba[..., :3].astype(np.float32) / 255.0) * alpha, alpha], axis=2)
    np.testing.assert_allclose(got, want, rtol=2e-6, atol=1e-7)    # from_rgba32 (common_device.cuh:771-798)
    tb.reload_network_from_json(IMAGE_CFG)
    tb.training_batch_size = 1 << 14
    tb.frame()
    first = tb.loss
    for _ in range(200):
        tb.frame()
    assert tb.loss < 0.2 * first, (first, tb.loss)

    # the reference's raw dump: int32 height, int32 width, RGBA halves (testbed_image.cu:420-437)
    half = want.astype(np.float16)
    with open(tmp_path / "picture.bin", "wb") as f:
        f.write(np.array([h, w], np.int32).tobytes()); f.write(half.tobytes())
    tb.load_training_data(str(tmp_path / "picture.bin"))
    got = tb.debug_tensor("image", torch.float32, h * w * 4).cpu().numpy().reshape(h, w, 4)
    np.testing.assert_array_equal(got, half.astype(np.float32))

    # the mode follows the data: a NeRF scene into the same testbed (testbed.cu:144-150), Testbed-level members stay
    _write_scene(tmp_path / "scene", views[:3])
    tb.load_training_data(str(tmp_path / "scene"))
    assert tb.mode == pyngp.TestbedMode.Nerf and tb.exposure == 0.5 and tb.nerf.training.dataset.n_images == 3
    tb.reload_network_from_json(SMALL_CFG)
    tb.train(BATCH)
    assert tb.training_step == 1
