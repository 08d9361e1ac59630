"""GPU parity of the SDF primitive's training step with caller-supplied ground truth (Testbed.override_sdf_training_data +
train_sdf, python_api.cu:69-99, testbed_sdf.cu:1289-1311): normalisation into the unit cube and the per-step permutation are
bit-exact against the oracle, the element-wise loss (MAPE of configs/sdf/base.json, and the other tcnn otypes) matches on the
GPU's own prediction, and a sphere SDF is learned."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SDF_CFG = {
    "loss": {"otype": "MAPE"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 10000, "decay_interval": 5000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-3, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.3819},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}


def _sphere_data(n, seed=3):
    """points in a mesh space [-2, 2]^3, signed distance to a sphere of radius 1.2 centred at (0.1, -0.2, 0.3)"""
    rng = np.random.default_rng(seed)
    p = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    d = (np.linalg.norm(p - np.array([0.1, -0.2, 0.3], np.float32), axis=1) - 1.2).astype(np.float32)
    return p, d


RAW_AABB = ((-2.0, -2.0, -2.0), (2.0, 2.0, 2.0))
MESH_SCALE = 4.0


@pytest.mark.parametrize("loss", ["MAPE", "L2", "RelativeL2", "L1", "SMAPE"])
def test_sdf_step_matches_oracle(ngp, ref, loss):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    tb = pyngp.Testbed(pyngp.TestbedMode.Sdf)
    n_data, n = 50_000, 1 << 13            # non power of two: the permutation's cycle walk is exercised
    p, d = _sphere_data(n_data)
    tb.reload_network_from_json({**SDF_CFG, "loss": {"otype": loss}})
    tb.override_sdf_training_data(p, d, RAW_AABB, MESH_SCALE)
    rp, rd = ref.sdf_normalize(p, d, RAW_AABB[0], RAW_AABB[1], MESH_SCALE)
    assert np.array_equal(tb.debug_tensor("sdf_positions", torch.float32, n_data * 3).cpu().numpy().reshape(n_data, 3), rp)
    assert np.array_equal(tb.debug_tensor("sdf_distances", torch.float32, n_data).cpu().numpy(), rd)
    assert rp.min() >= 0 and rp.max() <= 1
    for step in range(3):
        tb.train(n)
        pos = tb.debug_tensor("image_pos", torch.float32, n * 3).cpu().numpy().reshape(n, 3)
        tgt = tb.debug_tensor("image_target", torch.float32, n).cpu().numpy()
        sp, sd = ref.sdf_shuffle(rp, rd, n, step)         # seed = m_training_step before the step
        assert np.array_equal(pos, sp) and np.array_equal(tgt, sd)
        out = tb.debug_tensor("image_out", torch.float16, n * 16).cpu().numpy().reshape(n, 16)
        dl = tb.debug_tensor("image_dl", torch.float16, n * 16).cpu().numpy().reshape(n, 16)
        total, rdl = ref.tcnn_loss(loss, out, sd.reshape(n, 1), 1)
        np.testing.assert_allclose(dl.astype(np.float32), rdl.astype(np.float32), rtol=2e-3, atol=1e-8)
        assert not dl[:, 1:].any()
        if step == 0:
            assert abs(tb.loss - total) <= 1e-4 * max(total, 1e-6), (tb.loss, total)      # loss read back every 16th step
    assert tb.training_step == 3


def test_sdf_step_skipped_when_data_smaller_than_batch(ngp):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    tb = pyngp.Testbed(pyngp.TestbedMode.Sdf)
    tb.reload_network_from_json(SDF_CFG)
    with pytest.raises(RuntimeError):
        tb.train(256)                                       # no data at all
    p, d = _sphere_data(1000)
    tb.override_sdf_training_data(p, d, RAW_AABB, MESH_SCALE)
    before = tb.param_buffer(0).clone()
    tb.train(1024)                                          # testbed_sdf.cu:1293: size < batch -> no step
    assert tb.training_step == 0 and torch.equal(before, tb.param_buffer(0))
    with pytest.raises(RuntimeError):
        tb.override_sdf_training_data(p[:, :2], d)


def test_sdf_training_learns_a_sphere(ngp, ops):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    tb = pyngp.Testbed(pyngp.TestbedMode.Sdf)
    p, d = _sphere_data(1 << 18)
    tb.reload_network_from_json(SDF_CFG)
    tb.override_sdf_training_data(p, d, RAW_AABB, MESH_SCALE)
    net = ops.NetworkWithInputEncoding(SDF_CFG, 3, 1)
    assert net.n_params() == tb.n_params()
    q, qd = _sphere_data(1 << 12, seed=9)
    q = (q + 2.0) / 4.0; qd = qd / 4.0
    qx = torch.from_numpy(q.astype(np.float32)).cuda()

    def mean_abs_err():
        y = net.inference(tb.param_buffer(2), qx).float().cpu().numpy()[:, 0]
        return float(np.abs(y - qd).mean())

    e0 = mean_abs_err()
    for _ in range(400):
        tb.train(1 << 14)
    e1 = mean_abs_err()
    assert e1 < 0.01 and e1 < 0.2 * e0, (e0, e1)
