"""Host-side checks of the evaluation harness: the tonemap restatement (oracle/frame_output_ref.py) against the published
closed forms of the curves, the error metrics of instant-ngp_b200/evaluation.py against an independent restatement of scripts/common.py that
calls scipy.ndimage.convolve1d (the very routine the reference's SSIM calls), and the host-side BoundingBox of the binding."""
import importlib

import numpy as np
import pytest
import torch


@pytest.fixture(scope="module")
def fo():
    return importlib.import_module("oracle.frame_output_ref")


@pytest.fixture(scope="module")
def ev():
    return importlib.import_module("instant-ngp_b200.evaluation")


def test_tonemap_curves_match_published_forms(fo):
    x = np.linspace(0, 8, 401, dtype=np.float64)
    rgb = np.stack([x, x * 0.5, x * 0.25], axis=-1)
    # ACES (Narkowicz 2016) with the 0.6 pre-exposure folded in: f(0.6 x), f(v) = v (2.51 v + 0.03) / (v (2.43 v + 0.59) + 0.14)
    v = 0.6 * rgb
    np.testing.assert_allclose(fo.tonemap_curve(rgb, fo.ACES), v * (2.51 * v + 0.03) / (v * (2.43 * v + 0.59) + 0.14), rtol=2e-5, atol=1e-6)
    # Hable / Uncharted 2: f(2 x) / f(W), f(v) = (v (A v + C B) + D E) / (v (A v + B) + D F) - E / F, W = 11.2
    A, B, C, D, E, F = 0.15, 0.50, 0.10, 0.20, 0.02, 0.30
    f = lambda v: (v * (A * v + C * B) + D * E) / (v * (A * v + B) + D * F) - E / F
    np.testing.assert_allclose(fo.tonemap_curve(rgb, fo.HABLE), f(2 * rgb) / f(11.2), rtol=1e-4, atol=1e-6)
    # Reinhard on luminance
    Y = rgb @ np.array([0.2126, 0.7152, 0.0722])
    np.testing.assert_allclose(fo.tonemap_curve(rgb, fo.REINHARD), rgb / (Y + 1)[..., None], rtol=1e-5, atol=1e-7)
    assert fo.tonemap_curve(np.float32([[-1.0, 2.0, 3.0]]), fo.IDENTITY)[0, 0] == -1.0      # identity does not clamp; the curves do
    assert fo.tonemap_curve(np.float32([[-1.0, 2.0, 3.0]]), fo.ACES)[0, 0] == pytest.approx(0.0, abs=1e-6)


def test_tonemap_pass_background_exposure_and_colour_spaces(fo):
    rng = np.random.default_rng(0)
    a = rng.random((5, 7, 1)).astype(np.float32)
    acc = np.concatenate([rng.random((5, 7, 3)).astype(np.float32) * a, a], axis=-1)
    bg = [0.2, 0.4, 0.6, 1.0]
    out = fo.tonemap_pass(acc, 1.0, bg, fo.LINEAR, fo.LINEAR, fo.IDENTITY)
    want = (acc[..., :3] + fo.srgb_to_linear(np.float32(bg[:3])) * (1 - acc[..., 3:4])) * 2.0
    np.testing.assert_allclose(out[..., :3], want, rtol=1e-6)
    np.testing.assert_allclose(out[..., 3], 1.0, rtol=1e-6)
    # transparent background leaves the frame alone; sRGB output encodes it
    out = fo.tonemap_pass(acc, 0.0, [1, 1, 1, 0], fo.LINEAR, fo.SRGB, fo.IDENTITY)
    np.testing.assert_allclose(out[..., :3], fo.linear_to_srgb(acc[..., :3]), rtol=1e-6)
    np.testing.assert_array_equal(out[..., 3], acc[..., 3])
    # a frame accumulated in sRGB takes the background as it is given and is decoded before exposure
    out = fo.tonemap_pass(acc, 0.0, bg, fo.SRGB, fo.LINEAR, fo.IDENTITY)
    np.testing.assert_allclose(out[..., :3], fo.srgb_to_linear(acc[..., :3] + np.float32(bg[:3]) * (1 - acc[..., 3:4])), rtol=1e-5, atol=1e-7)


def test_accumulate_is_the_running_mean(fo):
    rng = np.random.default_rng(1)
    frames = rng.random((6, 4, 4, 4)).astype(np.float32)
    acc = np.zeros((4, 4, 4), np.float32)
    for i, f in enumerate(frames):
        acc = fo.accumulate(f, acc, i, fo.LINEAR)
    np.testing.assert_allclose(acc, frames.mean(axis=0), rtol=1e-5)
    acc = np.zeros((4, 4, 4), np.float32)
    for i, f in enumerate(frames):
        acc = fo.accumulate(f, acc, i, fo.SRGB)
    np.testing.assert_allclose(acc[..., :3], fo.linear_to_srgb(frames[..., :3]).mean(axis=0), rtol=1e-5)


def _common_py_restatement(metric, img, ref):
    """scripts/common.py:189-287 with scipy's convolve1d, written for this test only"""
    from scipy.ndimage import convolve1d
    img = np.where(np.isfinite(img), img, 0.0); img = np.maximum(img, 0.0)
    lum = lambda a: 0.2126 * a[:, :, 0] + 0.7152 * a[:, :, 1] + 0.0722 * a[:, :, 2]
    if metric == "SSIM":
        k = np.array([0.120078, 0.233881, 0.292082, 0.233881, 0.120078])
        blur = lambda a: convolve1d(convolve1d(a, k, axis=0), k, axis=1)
        a, b = lum(np.clip(img, 0, 1)), lum(np.clip(ref, 0, 1))
        mA, mB = blur(a), blur(b)
        sA, sB, sAB = blur(a * a) - mA ** 2, blur(b * b) - mB ** 2, blur(a * b) - mA * mB
        m = (2 * mA * mB + 1e-4) / (mA * mA + mB * mB + 1e-4) * (2 * sAB + 9e-4) / (sA + sB + 9e-4)
    else:
        l1, l2 = np.abs(img - ref), (img - ref) ** 2
        m = {"MAE": l1, "MAPE": l1 / (1e-2 + ref), "SMAPE": l1 / (1e-2 + (ref + img) / 2), "MSE": l2,
             "MScE": (np.clip(img, 0, 1) - np.clip(ref, 0, 1)) ** 2, "MRSE": l2 / (1e-2 + ref ** 2),
             "MRScE": (np.clip(img, 0, 100) - np.clip(ref, 0, 100)) ** 2 / (1e-2 + np.clip(ref, 0, 100) ** 2)}[metric]
    m = np.where(np.isfinite(m), m, 0.0)
    if m.ndim == 3:
        m = m.mean(axis=2)
    return float(m.mean())


@pytest.mark.parametrize("metric", ["MAE", "MAPE", "SMAPE", "MSE", "MScE", "MRSE", "MRScE", "SSIM"])
def test_error_metrics_match_the_reference_scripts(ev, metric):
    rng = np.random.default_rng(7)
    ref = rng.random((37, 53, 3)) * 1.5
    img = ref + rng.normal(0, 0.1, ref.shape)
    img[3, 4, 1] = np.nan; img[5, 6, 0] = np.inf
    want = _common_py_restatement(metric, img.copy(), ref)
    assert ev.compute_error(metric, img.copy(), ref) == pytest.approx(want, rel=1e-9)
    # torch input (float32, as frames come off the device) gives the same number to float32 precision
    got32 = ev.compute_error(metric, torch.from_numpy(img.astype(np.float32)), torch.from_numpy(ref.astype(np.float32)))
    assert got32 == pytest.approx(want, rel=2e-4)


def test_ssim_identity_trim_and_psnr(ev):
    rng = np.random.default_rng(3)
    a = rng.random((16, 16, 3))
    np.testing.assert_allclose(ev.SSIM(a, a), 1.0, rtol=1e-9)
    assert ev.compute_error("SSIM", a, 1 - a) < 0.5
    e = np.arange(1000000, dtype=np.float64)
    assert ev.trim(e) == pytest.approx(e[1:-1].mean())
    assert ev.compute_error("MtRSE", a, a) == 0.0
    assert ev.mse2psnr(0.01) == pytest.approx(20.0)
    x = np.linspace(0, 1, 11)
    np.testing.assert_allclose(ev.srgb_to_linear(ev.linear_to_srgb(x)), x, atol=1e-12)
    with pytest.raises(ValueError):
        ev.compute_error("nope", a, a)


def test_bounding_box_host_class():
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    b = pyngp.BoundingBox()
    assert b.is_empty()
    b.enlarge([0, 0, 0]); b.enlarge([1, 2, 3])
    assert not b.is_empty() and np.array_equal(b.diag(), [1, 2, 3]) and np.array_equal(b.center(), [0.5, 1, 1.5])
    assert b.contains([0.5, 1, 1]) and not b.contains([1.5, 1, 1])
    assert b.ray_intersect([-1, 1, 1], [1, 0, 0]) == (1.0, 2.0)
    assert b.ray_intersect([-1, 5, 1], [1, 0, 0])[0] == float(np.finfo(np.float32).max)
    assert b.distance([2, 2, 3]) == 1.0 and b.distance_sq([0.5, 1, 1]) == 0.0
    assert b.signed_distance([0.5, 1.0, 1.5]) == -0.5
    c = pyngp.BoundingBox([0.5, 0.5, 0.5], [4, 4, 4])
    i = b.intersection(c)
    assert np.array_equal(i.min, [0.5, 0.5, 0.5]) and np.array_equal(i.max, [1, 2, 3]) and b.intersects(c)
    assert not b.intersects(pyngp.BoundingBox([5, 5, 5], [6, 6, 6]))
    b.inflate(0.5)
    assert np.array_equal(b.min, [-0.5, -0.5, -0.5]) and len(b.get_vertices()) == 8
    np.testing.assert_allclose(b.relative_pos([0.5, 1, 1.5]), [0.5, 0.5, 0.5])
