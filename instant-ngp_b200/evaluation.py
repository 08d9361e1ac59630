"""Image-error half of the evaluation harness (reference: scripts/common.py:139-290, used by scripts/run.py:208-266): colour-space
conversion, per-pixel error maps (MAE / MAPE / SMAPE / MSE / MScE / MRSE / MtRSE / MRScE / SSIM) and their mean, `mse2psnr`.

Same names and argument meaning as the reference's helpers, so `run.py`'s evaluation loop reads unchanged.  The arithmetic runs in
torch on whatever device the images live on: frames rendered with `Testbed.render(..., out=<cuda tensor>)` and the device copies of the
training views are scored without a trip through host memory; numpy arrays are accepted and returned as numpy (float64 there, like the
reference's numpy code).  FLIP needs the third-party `flip` package of the reference's requirements.txt, which is not in this image."""
import numpy as np
import torch

# 5-tap binomial-ish blur of the reference's SSIM (common.py:190-194)
_SSIM_KERNEL = (0.120078, 0.233881, 0.292082, 0.233881, 0.120078)


def _t(x):
    """numpy / torch -> (torch tensor, was_numpy).  numpy input is computed in float64 on the host (what numpy itself would do)."""
    if isinstance(x, torch.Tensor):
        return x, False
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)), True


def _out(x, was_numpy):
    return x.cpu().numpy() if was_numpy else x


def mse2psnr(x):
    return -10.0 * np.log(x) / np.log(10.0)


def srgb_to_linear(img):
    t, n = _t(img)
    return _out(torch.where(t > 0.04045, ((t + 0.055) / 1.055).clamp_min(0) ** 2.4, t / 12.92), n)


def linear_to_srgb(img):
    t, n = _t(img)
    return _out(torch.where(t > 0.0031308, 1.055 * t.clamp_min(0) ** (1.0 / 2.4) - 0.055, 12.92 * t), n)


def luminance(a):
    return 0.2126 * a[:, :, 0] + 0.7152 * a[:, :, 1] + 0.0722 * a[:, :, 2]


def _blur(a):
    """separable 5-tap filter along both image axes with scipy.ndimage.convolve1d's default boundary ("reflect": the edge sample is
    repeated, d c b a | a b c d | d c b a), which is what common.py:191-194 calls"""
    k = torch.tensor(_SSIM_KERNEL, dtype=a.dtype, device=a.device)
    r = len(_SSIM_KERNEL) // 2
    for axis in (0, 1):
        a = a.movedim(axis, -1)
        padded = torch.cat([a[..., :r].flip(-1), a, a[..., -r:].flip(-1)], dim=-1)
        a = sum(k[j] * padded[..., j:j + a.shape[-1]] for j in range(2 * r + 1))      # symmetric kernel: convolution == correlation
        a = a.movedim(-1, axis)
    return a


def SSIM(a, b):
    """per-pixel structural similarity of the luminance channels (common.py:189-207)"""
    ta, n = _t(a); tb, _ = _t(b)
    la, lb = luminance(ta), luminance(tb)
    mA, mB = _blur(la), _blur(lb)
    sA, sB, sAB = _blur(la * la) - mA ** 2, _blur(lb * lb) - mB ** 2, _blur(la * lb) - mA * mB
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    p1 = (2.0 * mA * mB + c1) / (mA * mA + mB * mB + c1)
    p2 = (2.0 * sAB + c2) / (sA + sB + c2)
    return _out(p1 * p2, n)


def L1(img, ref):
    return abs(img - ref)


def APE(img, ref):
    return L1(img, ref) / (1e-2 + ref)


def SAPE(img, ref):
    return L1(img, ref) / (1e-2 + (ref + img) / 2.)


def L2(img, ref):
    return (img - ref) ** 2


def RSE(img, ref):
    return L2(img, ref) / (1e-2 + ref ** 2)


def trim(error, skip=0.000001):
    t, n = _t(error)
    flat = torch.sort(t.flatten()).values
    size = flat.numel()
    s = int(skip * size)
    m = flat[s:size - s].mean()
    return float(m) if n else m


def compute_error_img(metric, img, ref):
    """error map of `img` against `ref` ([H,W,C], linear colours unless the caller converted them); non-finite pixels of img count as 0
    and negative ones are clamped, as in common.py:236-238 (which does so in place; this function leaves its arguments alone)"""
    ti, n = _t(img); tr, _ = _t(ref)
    tr = tr.to(ti.dtype)
    ti = torch.where(torch.isfinite(ti), ti, torch.zeros_like(ti)).clamp_min(0.0)
    if metric == "MAE":
        e = L1(ti, tr)
    elif metric == "MAPE":
        e = APE(ti, tr)
    elif metric == "SMAPE":
        e = SAPE(ti, tr)
    elif metric == "MSE":
        e = L2(ti, tr)
    elif metric == "MScE":
        e = L2(ti.clamp(0.0, 1.0), tr.clamp(0.0, 1.0))
    elif metric == "MRSE":
        e = RSE(ti, tr)
    elif metric == "MtRSE":
        e = trim(RSE(ti, tr))
        return float(e) if n else e
    elif metric == "MRScE":
        e = RSE(ti.clamp(0, 100), tr.clamp(0, 100))
    elif metric == "SSIM":
        e = SSIM(ti.clamp(0.0, 1.0), tr.clamp(0.0, 1.0))
    elif metric in ("FLIP", "\\FLIP"):
        raise ValueError("FLIP needs the third-party `flip` package (requirements.txt of the reference), which this image does not have")
    else:
        raise ValueError(f"Unknown metric: {metric}.")
    return _out(e, n)


def compute_error(metric, img, ref):
    """mean of the error map (channels averaged first), non-finite entries of the map counted as 0 (common.py:280-287); a Python float"""
    m = compute_error_img(metric, img, ref)
    if not isinstance(m, (torch.Tensor, np.ndarray)):
        return float(m)
    t = m if isinstance(m, torch.Tensor) else torch.from_numpy(m)
    t = torch.where(torch.isfinite(t), t, torch.zeros_like(t))
    if t.dim() == 3:
        t = t.mean(dim=2)
    return float(t.mean())
