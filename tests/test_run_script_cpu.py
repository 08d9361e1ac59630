"""What scripts/run.py asks of the binding before any kernel runs (run.py:88-170): a Testbed without a mode, root_dir, the mode picked from the
scene (mode_from_scene, common.cu:144-159), options set before the mode is known, the GUI-only calls, configs with comments.  No device needed."""
import importlib
import json

import pytest


@pytest.fixture(scope="module")
def pyngp(ngp):
    return importlib.import_module("instant-ngp_b200.pyngp")


def test_mode_from_scene(pyngp, tmp_path):
    (tmp_path / "scene").mkdir()
    for name in ("transforms.JSON", "bunny.obj", "bunny.STL", "cloud.nvdb", "albert.exr", "photo.jpg", "dump.bin", "noext"):
        (tmp_path / name).write_bytes(b"")
    M = pyngp.TestbedMode
    expect = {"scene": M.Nerf, "scene/": M.Nerf, "transforms.JSON": M.Nerf, "bunny.obj": M.Sdf, "bunny.STL": M.Sdf, "cloud.nvdb": M.Volume,
              "albert.exr": M.Image, "photo.jpg": M.Image, "dump.bin": M.Image, "noext": M.Image, "missing.json": M.None_}
    for name, mode in expect.items():
        assert pyngp.mode_from_scene(tmp_path / name) == mode, name


def test_testbed_without_a_mode(pyngp, tmp_path):
    tb = pyngp.Testbed()
    assert tb.mode == pyngp.TestbedMode.None_
    tb.root_dir = str(tmp_path)
    assert tb.root_dir == str(tmp_path)
    assert tb.want_repl() is False
    # members of Testbed itself are kept for the mode to come (run.py sets them after loading, a notebook may set them before)
    tb.exposure = 1.5
    tb.background_color = [0.0, 0.0, 0.0, 1.0]
    tb.color_space = pyngp.ColorSpace.SRGB
    tb.shall_train = False
    assert tb.exposure == 1.5 and list(tb.background_color) == [0.0, 0.0, 0.0, 1.0] and tb.color_space == pyngp.ColorSpace.SRGB
    tb.nerf.sharpen = 0.0
    assert tb.nerf.sharpen == 0.0
    with pytest.raises(RuntimeError, match="sharpen"):
        tb.nerf.sharpen = 0.5
    # mode-specific state does not exist yet
    with pytest.raises(RuntimeError, match="no mode yet"):
        tb.nerf.training.near_distance = 0.2
    with pytest.raises(RuntimeError, match="no mode yet"):
        tb.train(256)
    with pytest.raises(RuntimeError, match="no mode yet"):
        tb.render(8, 8, 1, True)
    # a network config named before the mode is known is only remembered (testbed.cu:288-292)
    tb.reload_network_from_file("small.json")
    assert tb._requested_config == "small.json"
    with pytest.raises(RuntimeError, match="does not exist"):
        tb.load_training_data(tmp_path / "nothing_here")
    with pytest.raises(RuntimeError, match="does not exist"):
        tb.load_file(tmp_path / "nothing.ingp")
    (tmp_path / "cloud.nvdb").write_bytes(b"")
    with pytest.raises(RuntimeError, match="Volume"):
        tb.load_training_data(tmp_path / "cloud.nvdb")
    assert tb.mode == pyngp.TestbedMode.None_


def test_gui_only_calls_say_so(pyngp):
    tb = pyngp.Testbed()
    for call, what in ((lambda: tb.init_window(1920, 1080, second_window=False), "GUI"), (tb.init_vr, "OpenXR"),
                       (lambda: tb.load_camera_path("base_cam.json"), "camera paths"),
                       (lambda: tb.compute_and_save_marching_cubes_mesh("m.obj", [256, 256, 256]), "marching cubes")):
        with pytest.raises(RuntimeError, match=what):
            call()


def test_json_with_comments(pyngp):
    text = '{\n\t"optimizer": {\n\t\t// "otype": "Ema",\n\t\t"otype": "Adam", /* block\n comment */ "url": "http://x//y", "q": "a\\"//b"\n\t}\n}'
    assert pyngp._loads_with_comments(text) == {"optimizer": {"otype": "Adam", "url": "http://x//y", "q": 'a"//b'}}


def test_builtin_base_configs(pyngp):
    """the defaults a testbed trains with when no config was given: the hyper-parameters of configs/<mode>/base.json"""
    nerf, sdf, image = (pyngp.builtin_base_config(m) for m in (pyngp.TestbedMode.Nerf, pyngp.TestbedMode.Sdf, pyngp.TestbedMode.Image))
    assert nerf["loss"]["otype"] == "Huber" and sdf["loss"]["otype"] == "MAPE" and image["loss"]["otype"] == "L2"
    assert (nerf["encoding"]["n_levels"], nerf["encoding"]["n_features_per_level"], nerf["encoding"]["log2_hashmap_size"]) == (4, 4, 19)
    assert (image["encoding"]["n_levels"], image["encoding"]["n_features_per_level"], image["encoding"]["log2_hashmap_size"]) == (16, 2, 24)
    assert nerf["optimizer"]["otype"] == "Ema" and nerf["optimizer"]["nested"]["nested"]["learning_rate"] == 1e-2
    assert image["optimizer"]["otype"] == "ExponentialDecay" and sdf["optimizer"]["nested"]["nested"]["learning_rate"] == 1e-4
    assert nerf["network"]["n_hidden_layers"] == 1 and nerf["rgb_network"]["n_hidden_layers"] == 2
    json.dumps(nerf)
