"""pyngp.Testbed.load_training_data (python_api.cu:398): a synthetic scene written to disk as transforms.json + PNG files in the
NeRF convention is loaded through the library's loader and trains like the same views handed over in memory."""
import importlib
import json

import msgpack
import numpy as np
import pytest

from tests.test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG

pytestmark = pytest.mark.gpu


def _ngp_to_nerf(x, scale=0.33, offset=0.5):
    """inverse of nerf_matrix_to_ngp (nerf_loader.h:142-158 ngp_matrix_to_nerf) on a column-major mat4x3 -> 4x4 rows"""
    m = np.asarray(x, np.float64).reshape(4, 3).T          # 3 rows x 4 columns, NGP
    m = m[[2, 0, 1], :]                                    # rows xyz -> yzx undone
    m[:, 1] *= -1; m[:, 2] *= -1
    m[:, 3] = (m[:, 3] - offset) / scale
    return np.vstack([m, [0, 0, 0, 1]]).tolist()


def test_load_training_data_from_disk(ngp, tmp_path):
    import cv2
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    frames = []
    (tmp_path / "train").mkdir()
    for i, v in enumerate(views):
        rgba = v["pixels"].numpy()
        cv2.imwrite(str(tmp_path / "train" / f"r_{i}.png"), np.ascontiguousarray(rgba[..., [2, 1, 0, 3]]))
        frames.append({"file_path": f"./train/r_{i}", "transform_matrix": _ngp_to_nerf(v["xform"])})
    frames.reverse()                                       # the loader restores natural order
    (tmp_path / "transforms.json").write_text(json.dumps({"camera_angle_x": scene.CAMERA_ANGLE_X, "aabb_scale": 1, "frames": frames}))

    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.load_training_data(tmp_path / "transforms.json")
    tb.reload_network_from_json(SMALL_CFG)
    ref = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    ref.set_nerf_dataset(views, aabb_scale=1)
    ref.reload_network_from_json(SMALL_CFG)
    # what the loader produced, read back through a snapshot: same order, poses and focal length as the in-memory views
    tb.save_snapshot(tmp_path / "s.msgpack")
    ds = msgpack.unpackb((tmp_path / "s.msgpack").read_bytes(), raw=False)["snapshot"]["nerf"]["dataset"]
    assert ds["n_images"] == N_VIEWS and ds["aabb_scale"] == 1
    for i, v in enumerate(views):
        got = np.asarray(ds["xforms"][i]["start"], np.float32).T.reshape(-1)
        np.testing.assert_allclose(got, np.asarray(v["xform"], np.float32), atol=2e-6)
        np.testing.assert_allclose(ds["metadata"][i]["focal_length"], v["focal_length"], rtol=1e-6)
        assert ds["metadata"][i]["resolution"] == [RES, RES]
    losses = {}
    for name, t in (("disk", tb), ("memory", ref)):
        for _ in range(49):
            t.train(BATCH)
        losses[name] = t.loss
    assert np.isfinite(losses["disk"]) and abs(losses["disk"] - losses["memory"]) < 0.2 * losses["memory"], losses
    with pytest.raises(RuntimeError, match="does not exist"):
        tb.load_training_data(tmp_path / "missing.json")
