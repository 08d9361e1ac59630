"""Generates tests/golden/rotvec.npz from the reference's OWN rotmat(vec3) / rotvec(mat3) (common_device.cuh:698-722, through the mount's GLM)
compiled into oracle/_ref/libngp_refshim.so (`make -C oracle ref`; oracle/ref_shim/shim.cpp).  Run where the mount exists:

    make -C oracle ref && python tests/golden/make_golden_rotvec.py

Angle-axis vectors from 1e-4 rad (one Adam step of the rotation optimiser) to beyond pi/2, the zero vector, and the composition
rotvec(rotmat(-step) * rotmat(variable)) that RotationAdamOptimizer::step evaluates (adam_optimizer.h:246)."""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
shim = C.CDLL(os.path.join(HERE, "..", "..", "oracle", "_ref", "libngp_refshim.so"))
vp = C.c_void_p
shim.shim_rotmat.argtypes = [vp, vp]; shim.shim_rotvec.argtypes = [vp, vp]
P = lambda a: a.ctypes.data_as(vp)
rng = np.random.default_rng(20261106)
N = 4000

mag = np.concatenate([[0.0], 10.0 ** rng.uniform(-4, 0.45, N - 1)]).astype(np.float32)      # up to ~2.8 rad
axis = rng.normal(size=(N, 3)); axis /= np.linalg.norm(axis, axis=1, keepdims=True)
vec = (axis * mag[:, None]).astype(np.float32)
mat = np.zeros((N, 9), np.float32); back = np.zeros((N, 3), np.float32)
for k in range(N):
    shim.shim_rotmat(P(vec[k]), P(mat[k]))
    shim.shim_rotvec(P(mat[k]), P(back[k]))
# the optimiser's composition: a step of ~1e-3 rad against a variable of up to 0.1 rad
step = (rng.normal(size=(N, 3)) * 1e-3).astype(np.float32); var = (rng.normal(size=(N, 3)) * 10.0 ** rng.uniform(-4, -1, (N, 1))).astype(np.float32)
composed = np.zeros((N, 3), np.float32)
for k in range(N):
    a = np.zeros(9, np.float32); b = np.zeros(9, np.float32)
    neg = (-step[k]).astype(np.float32)
    shim.shim_rotmat(P(neg), P(a)); shim.shim_rotmat(P(var[k]), P(b))
    A, B = a.reshape(3, 3).T, b.reshape(3, 3).T                         # column major -> rows
    prod = np.zeros((3, 3), np.float32)                                  # glm mat3 * mat3 in float32: sum over k of A[:, k] * B[k, c], left to right
    for c in range(3):
        for r in range(3):
            prod[r, c] = np.float32(np.float32(np.float32(A[r, 0] * B[0, c]) + np.float32(A[r, 1] * B[1, c])) + np.float32(A[r, 2] * B[2, c]))
    m = np.ascontiguousarray(prod.T.reshape(-1))
    shim.shim_rotvec(P(m), P(composed[k]))
np.savez_compressed(os.path.join(HERE, "rotvec.npz"), vec=vec, mat=mat, back=back, step=step, var=var, composed=composed)
rel = np.linalg.norm(back[1:] - vec[1:], axis=1) / np.linalg.norm(vec[1:], axis=1)
print("wrote", N, "cases; rotvec(rotmat(v)) vs v: max relative deviation", float(rel.max()), "at |v| =", float(mag[1:][rel.argmax()]))
