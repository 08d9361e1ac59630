"""GPU parity of the image primitive (BASELINE config 2 shape): sample generation + target lookup bit-exact vs the oracle,
L2 loss gradients, convergence (compute_image_mse)."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

IMG_CFG = {
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.5},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}


def _image(w=160, h=128):
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.zeros((h, w, 4), np.float32)
    img[..., 0] = 0.5 + 0.5 * np.sin(x / 9.0) * np.cos(y / 13.0)
    img[..., 1] = ((x // 16 + y // 16) % 2) * 0.8 + 0.1
    img[..., 2] = np.clip((x + y) / (w + h), 0, 1)
    img[..., 3] = 1.0
    return img


@pytest.mark.parametrize("snap", [True, False])
def test_image_batch_matches_oracle(ngp, ref, snap):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    tb = pyngp.Testbed(pyngp.TestbedMode.Image)
    img = _image()
    tb.set_image(img)
    tb.reload_network_from_json(IMG_CFG)
    tb.image_snap_to_pixel_centers = snap
    n = 1 << 14
    tb.train(n)
    pos = tb.debug_tensor("image_pos", torch.float32, n * 2).cpu().numpy().reshape(n, 2)
    tgt = tb.debug_tensor("image_target", torch.float32, n * 3).cpu().numpy().reshape(n, 3)
    rpos = np.zeros((n, 2), np.float32)
    ref.lib().ref_image_positions(ref.u64(1337), ref.u64(1), ref.u32(n), ref.i32(1), ref.P(rpos))   # m_rng: one next_uint consumed at reset
    rtgt = np.zeros((n, 3), np.float32)
    ref.lib().ref_image_eval(ref.P(img), ref.i32(img.shape[1]), ref.i32(img.shape[0]), ref.u32(n), ref.P(rpos), ref.P(rtgt), ref.i32(int(snap)), ref.i32(0))
    assert np.array_equal(pos, rpos)
    np.testing.assert_allclose(tgt, rtgt, rtol=1e-5, atol=1e-6)
    # L2 loss gradient on the GPU's own prediction
    out = tb.debug_tensor("image_out", torch.float16, n * 16).cpu().numpy().reshape(n, 16)
    dl = tb.debug_tensor("image_dl", torch.float16, n * 16).cpu().numpy().reshape(n, 16)
    rdl = np.zeros((n, 16), np.float16)
    ref.lib().ref_l2_loss(ref.P(out), 16, ref.P(rtgt), n, 128.0, ref.P(rdl))
    np.testing.assert_allclose(dl.astype(np.float32), rdl.astype(np.float32), rtol=2e-3, atol=1e-8)


@pytest.mark.parametrize("mode", ["Halton", "Sobol"])
def test_image_low_discrepancy_samplers_match_oracle(ngp, ref, mode):
    """image.random_mode = Halton | Sobol (testbed_image.cu:41-53,227-231): positions of two consecutive steps (the sequences are indexed by
    batch_size * training_step) bit-exact against the restated generators, targets as for the random modes."""
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    tb = pyngp.Testbed(pyngp.TestbedMode.Image)
    img = _image()
    tb.set_image(img)
    tb.reload_network_from_json(IMG_CFG)
    tb.image.random_mode = getattr(pyngp.RandomMode, mode)
    assert tb.image.random_mode == getattr(pyngp.RandomMode, mode)
    n = 1 << 12
    for step in range(2):
        tb.train(n)
        pos = tb.debug_tensor("image_pos", torch.float32, n * 2).cpu().numpy().reshape(n, 2)
        tgt = tb.debug_tensor("image_target", torch.float32, n * 3).cpu().numpy().reshape(n, 3)
        rpos = np.zeros((n, 2), np.float32)
        ref.lib().ref_image_positions_ld(ref.i32(int(getattr(pyngp.RandomMode, mode))), ref.u64(n * step), ref.u32(1337), ref.u32(n), ref.P(rpos))
        raw = rpos.copy()
        rtgt = np.zeros((n, 3), np.float32)
        ref.lib().ref_image_eval(ref.P(img), ref.i32(img.shape[1]), ref.i32(img.shape[0]), ref.u32(n), ref.P(rpos), ref.P(rtgt), ref.i32(1), ref.i32(0))
        assert np.array_equal(pos, rpos)
        np.testing.assert_allclose(tgt, rtgt, rtol=1e-5, atol=1e-6)
        assert raw.min() >= 0.0 and raw.max() < 1.0
        # low discrepancy: a 16 x 16 grid of cells receives n / 256 = 16 points each, give or take (a random draw would spread far wider)
        counts = np.histogram2d(raw[:, 0], raw[:, 1], bins=16, range=[[0, 1], [0, 1]])[0]
        assert counts.min() >= 8 and counts.max() <= 24, (counts.min(), counts.max())


def test_image_training_converges(ngp):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    tb = pyngp.Testbed(pyngp.TestbedMode.Image)
    tb.set_image(_image())
    tb.reload_network_from_json(IMG_CFG)
    mse0 = tb.compute_image_mse()
    for _ in range(300):
        tb.train(1 << 14)
    mse1 = tb.compute_image_mse()
    assert mse1 < 0.05 * mse0 and mse1 < 2e-3, (mse0, mse1)
    assert tb.training_step == 300 and tb.loss > 0


# ---------------------------------------------------------------------------------------------------------------------
# BASELINE config 2: data/image/albert.exr (staged by tools/stage_data.py into baseline/_data/, 1024x1024 RGBA float32),
# 2-D HashGrid 16x2 T = 2^24 + 2x64 FullyFusedMLP (configs/image/base.json), 2^18 pixels per batch
# ---------------------------------------------------------------------------------------------------------------------
ALBERT = __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))),
                                   "baseline", "_data", "albert", "albert_rgba_f32.npy")
ALBERT_CFG = {**IMG_CFG, "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24, "base_resolution": 16,
                                      "per_level_scale": 2.0}}


@pytest.mark.skipif(not __import__("os").path.exists(ALBERT), reason="albert.exr not staged (python tools/stage_data.py in the build container)")
def test_config2_albert_t24_batch_2pow18(ngp, ref):
    """one training step at the configuration BASELINE.json names: positions bit-exact, bilinear / nearest targets, L2 gradients vs the oracle
    on the real image; then 60 steps must cut the full-image MSE (compute_image_mse over all 2^20 pixels) by 10x"""
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    img = np.load(ALBERT)
    assert img.shape == (1024, 1024, 4) and img.dtype == np.float32
    tb = pyngp.Testbed(pyngp.TestbedMode.Image)
    tb.set_image(img)
    tb.reload_network_from_json(ALBERT_CFG)
    assert tb.n_params() > 250_000_000           # 2-D, 16 levels, T = 2^24: ~280 M parameters
    n = 1 << 18
    mse0 = tb.compute_image_mse()
    tb.train(n)
    pos = tb.debug_tensor("image_pos", torch.float32, n * 2).cpu().numpy().reshape(n, 2)
    tgt = tb.debug_tensor("image_target", torch.float32, n * 3).cpu().numpy().reshape(n, 3)
    rpos = np.zeros((n, 2), np.float32)
    ref.lib().ref_image_positions(ref.u64(1337), ref.u64(1), ref.u32(n), ref.i32(1), ref.P(rpos))
    rtgt = np.zeros((n, 3), np.float32)
    ref.lib().ref_image_eval(ref.P(img), ref.i32(1024), ref.i32(1024), ref.u32(n), ref.P(rpos), ref.P(rtgt), ref.i32(1), ref.i32(0))
    assert np.array_equal(pos, rpos)
    np.testing.assert_allclose(tgt, rtgt, rtol=1e-5, atol=1e-6)
    out = tb.debug_tensor("image_out", torch.float16, n * 16).cpu().numpy().reshape(n, 16)
    dl = tb.debug_tensor("image_dl", torch.float16, n * 16).cpu().numpy().reshape(n, 16)
    rdl = np.zeros((n, 16), np.float16)
    ref.lib().ref_l2_loss(ref.P(out), 16, ref.P(rtgt), n, 128.0, ref.P(rdl))
    np.testing.assert_allclose(dl.astype(np.float32), rdl.astype(np.float32), rtol=2e-3, atol=1e-8)
    for _ in range(60):
        tb.train(n)
    mse1 = tb.compute_image_mse()
    assert mse1 < 0.1 * mse0, (mse0, mse1)
