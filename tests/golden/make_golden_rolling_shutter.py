"""Generates tests/golden/rolling_shutter.npz from the reference's OWN get_xform_given_rolling_shutter (common_device.cuh:401-408, evaluated
through the mount's GLM) compiled into oracle/_ref/libngp_refshim.so (`make -C oracle ref`; oracle/ref_shim/shim.cpp).  Run where the mount exists:

    make -C oracle ref && python tests/golden/make_golden_rolling_shutter.py

Seeded poses: identical start / end (what every ordinary data set has: the quaternion round trip alone), nearly identical (glm's linear
branch of slerp), and clearly different ones (the acos / sin branch)."""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
shim = C.CDLL(os.path.join(HERE, "..", "..", "oracle", "_ref", "libngp_refshim.so"))
vp, f32 = C.c_void_p, C.c_float
shim.shim_get_xform_given_rolling_shutter.argtypes = [vp, vp, vp, f32, f32, f32, vp]
P = lambda a: a.ctypes.data_as(vp)
rng = np.random.default_rng(20261105)
N = 3000


def pose():
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    m = np.zeros(12, np.float32)
    m[:9] = q.T.astype(np.float32).reshape(-1)          # column major: columns of the rotation
    m[9:] = rng.uniform(-1, 2, 3).astype(np.float32)
    return m


def perturb(m, amount):
    w = rng.normal(size=3) * amount
    k = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    r = np.eye(3) + k + k @ k / 2
    u, _, vt = np.linalg.svd(r)
    rot = (u @ vt) @ m[:9].reshape(3, 3).T.astype(np.float64)
    out = m.copy(); out[:9] = rot.T.astype(np.float32).reshape(-1); out[9:] += rng.normal(size=3).astype(np.float32) * np.float32(amount)
    return out


start = np.stack([pose() for _ in range(N)])
end = start.copy()
end[N // 3: 2 * N // 3] = np.stack([perturb(m, 1e-5) for m in start[N // 3: 2 * N // 3]])
end[2 * N // 3:] = np.stack([perturb(m, 0.3) for m in start[2 * N // 3:]])
rs = rng.uniform(-0.2, 0.6, (N, 4)).astype(np.float32); rs[: N // 6] = 0
uvt = rng.uniform(0, 1, (N, 3)).astype(np.float32)
out = np.zeros((N, 12), np.float32)
for k in range(N):
    shim.shim_get_xform_given_rolling_shutter(P(start[k]), P(end[k]), P(rs[k]), float(uvt[k, 0]), float(uvt[k, 1]), float(uvt[k, 2]), P(out[k]))
np.savez_compressed(os.path.join(HERE, "rolling_shutter.npz"), start=start, end=end, rolling_shutter=rs, uvt=uvt, xform=out)
print("wrote", N, "cases; max |xform - start| over the identical-pose third:", float(np.abs(out[: N // 3, :9] - start[: N // 3, :9]).max()))
