"""Pins the oracle's in-mount restatements against the reference's OWN code: tests/golden/reference_helpers.npz was produced
by compiling include/neural-graphics-primitives/{random_val,bounding_box,common_device}.cuh and the helper functions of
src/testbed_nerf.cu with g++ (oracle/ref_shim, tests/golden/make_golden.py).  Integer results must be identical; float
results must be bit-identical too (both sides are IEEE evaluations of the same expressions) except where noted."""
import ctypes as C
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_helpers.npz"))


@pytest.fixture(scope="module")
def L(ref):
    """the oracle in math mode 0: libm logf/expf, i.e. what the reference's own helper code compiled with g++ evaluates — this
    pins the FORMULAS; mode 1 (the default everywhere else) swaps in the shared explicitly rounded logf/expf that the CUDA
    kernels compile too (include/ngp_b200_math.h), checked against libm in test_shared_logf_expf_vs_libm"""
    l = ref.lib()
    l.ref_set_math_mode(0)
    yield from _bind(l)
    l.ref_set_math_mode(1)


def _bind(l):
    f32, u32, i32, vp = C.c_float, C.c_uint32, C.c_int, C.c_void_p
    for name, res, args in [("ref_advance_to_next_voxel", f32, [f32, f32, vp, vp, u32]), ("ref_warp_dt", f32, [f32]), ("ref_unwarp_dt", f32, [f32]),
                            ("ref_network_to_rgb", f32, [f32, i32]), ("ref_network_to_rgb_derivative", f32, [f32, i32]),
                            ("ref_network_to_density", f32, [f32, i32]), ("ref_network_to_density_derivative", f32, [f32, i32]),
                            ("ref_mip_from_pos", u32, [vp, u32]), ("ref_mip_from_dt", u32, [f32, vp, u32]), ("ref_cascaded_grid_idx_at", u32, [vp, u32]),
                            ("ref_density_grid_occupied_at", i32, [vp, vp, u32]), ("ref_aabb_contains", i32, [vp, vp])]:
        getattr(l, name).restype = res; getattr(l, name).argtypes = args
    yield l


def _same(a, b):
    a = np.asarray(a); b = np.asarray(b)
    return np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b)


def test_low_discrepancy_sequence(ref, L):
    got = np.array([L.ref_ld_random_val(int(a), int(b)) for a, b in zip(G["ld_index"], G["ld_seed"])], np.float32)
    assert _same(got, G["ld_val"])
    v2 = np.zeros_like(G["ld_val_2d"])
    for k in range(len(v2)):
        L.ref_ld_random_val_2d(C.c_uint32(int(G["ld_index"][k])), C.c_uint32(int(G["ld_seed"][k])), ref.P(v2[k]))
    assert _same(v2, G["ld_val_2d"])


def test_colour_conversions(L):
    x = G["colour_x"]
    assert _same(np.array([L.ref_srgb_to_linear(float(a)) for a in x], np.float32), G["srgb_to_linear"])
    assert _same(np.array([L.ref_linear_to_srgb(float(a)) for a in x], np.float32), G["linear_to_srgb"])


@pytest.mark.parametrize("tag,cone", [("c0", 0.0), ("c256", 1.0 / 256.0)])
def test_exponential_stepping(L, tag, cone):
    t = G["step_t"]
    n = np.array([L.ref_to_stepping_space(float(a), cone) for a in t], np.float32)
    assert _same(n, G[f"to_stepping_{tag}"])
    assert _same(np.array([L.ref_from_stepping_space(float(a), cone) for a in n], np.float32), G[f"from_stepping_{tag}"])
    assert _same(np.array([L.ref_calc_dt(float(a), cone) for a in t], np.float32), G[f"calc_dt_{tag}"])
    assert _same(np.array([L.ref_advance_n_steps(float(a), cone, 0.5) for a in t], np.float32), G[f"advance_half_{tag}"])


def test_occupancy_lookups_and_voxel_skip(ref, L):
    pos, d, tt = G["occ_pos"], G["occ_dir"], G["occ_t"]
    bitfield = np.random.default_rng(int(G["bitfield_seed"][0])).integers(0, 256, 128 ** 3 * 8 // 8, dtype=np.uint16).astype(np.uint8)
    for mc in (0, 3, 7):
        assert np.array_equal(np.array([L.ref_mip_from_pos(ref.P(p), mc) for p in pos], np.uint32), G[f"mip_from_pos_{mc}"])
        assert np.array_equal(np.array([L.ref_mip_from_dt(float(dt), ref.P(p), mc) for p, dt in zip(pos, G["calc_dt_c256"])], np.uint32), G[f"mip_from_dt_{mc}"])
    mips = G["mip_from_pos_7"]
    assert np.array_equal(np.array([L.ref_cascaded_grid_idx_at(ref.P(p), int(m)) for p, m in zip(pos, mips)], np.uint32), G["grid_idx"])
    occ = np.array([L.ref_density_grid_occupied_at(ref.P(p), ref.P(bitfield), int(m)) for p, m in zip(pos, mips)], np.int32)
    assert np.array_equal(occ != 0, G["occupied"] != 0) and 0.2 < (occ != 0).mean() < 0.8
    for tag, cone in (("c0", 0.0), ("c256", 1.0 / 256.0)):
        nv = np.array([L.ref_advance_to_next_voxel(float(a), cone, ref.P(p), ref.P(q), int(m)) for a, p, q, m in zip(tt, pos, d, mips)], np.float32)
        assert _same(nv, G[f"next_voxel_{tag}"])


def test_bounding_box(ref, L):
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    out = np.zeros((len(G["aabb_o"]), 2), np.float32)
    for k in range(len(out)):
        L.ref_aabb_ray_intersect(C.byref(aabb), ref.P(G["aabb_o"][k]), ref.P(G["occ_dir"][k]), ref.P(out[k]))
    assert _same(out, G["aabb_isect"])
    assert np.array_equal(np.array([L.ref_aabb_contains(C.byref(aabb), ref.P(p)) for p in G["occ_pos"]], np.int32) != 0, G["aabb_contains"] != 0)


def test_dt_warp_activations_and_losses(ref, L):
    dt = G["dt"]
    w = np.array([L.ref_warp_dt(float(a)) for a in dt], np.float32)
    assert _same(w, G["warp_dt"])
    assert _same(np.array([L.ref_unwarp_dt(float(a)) for a in w], np.float32), G["unwarp_dt"])
    for act in range(4):
        for fn in ("network_to_rgb", "network_to_rgb_derivative", "network_to_density", "network_to_density_derivative"):
            got = np.array([getattr(L, "ref_" + fn)(float(a), act) for a in G["act_x"]], np.float32)
            assert _same(got, G[f"{fn}_{act}"]), (fn, act)
    for lt in range(7):
        lg = np.zeros_like(G[f"loss_{lt}"])
        for k in range(len(lg)):
            L.ref_loss_and_gradient(ref.P(G["loss_target"][k]), ref.P(G["loss_pred"][k]), lt, ref.P(lg[k]))
        assert _same(lg, G[f"loss_{lt}"]), lt


def _ulp_diff(a, b):
    ia = a.view(np.int32).astype(np.int64); ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia); ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    return np.abs(ia - ib)


def test_shared_logf_expf_vs_libm(ref):
    """include/ngp_b200_math.h (compiled into the oracle AND the CUDA library from the same source): within 2 ulp of the
    correctly rounded result over the ranges the marcher and the compositor feed it, exact IEEE conventions at the edges"""
    l = ref.lib()
    l.ref_shared_logf.restype = C.c_float; l.ref_shared_logf.argtypes = [C.c_float]
    l.ref_shared_expf.restype = C.c_float; l.ref_shared_expf.argtypes = [C.c_float]
    rng = np.random.default_rng(7)
    x = np.concatenate([np.exp(rng.uniform(-80, 80, 20000)), rng.uniform(0.5, 2.0, 20000), rng.uniform(1e-3, 64.0, 20000),
                        np.array([1.0, 1.0 + 1.0 / 256.0, 1.73205080757 / 1024, 1e-38, 1e-42, 3.4e38])]).astype(np.float32)
    got = np.array([l.ref_shared_logf(float(v)) for v in x], np.float32)
    want = np.log(x.astype(np.float64)).astype(np.float32)
    assert _ulp_diff(got, want).max() <= 2
    y = np.concatenate([rng.uniform(-100, 88, 40000), rng.uniform(-1, 1, 20000), np.array([0.0, -0.0, 88.7, -103.9, -87.5])]).astype(np.float32)
    got = np.array([l.ref_shared_expf(float(v)) for v in y], np.float32)
    want = np.exp(y.astype(np.float64)).astype(np.float32)
    normal = want > 1.2e-38
    assert _ulp_diff(got[normal], want[normal]).max() <= 2
    assert np.abs(got[~normal].astype(np.float64) - want[~normal].astype(np.float64)).max() <= 2 * 1.4e-45 + 1e-44
    assert l.ref_shared_expf(100.0) == np.inf and l.ref_shared_expf(-200.0) == 0.0 and l.ref_shared_logf(0.0) == -np.inf
    assert np.isnan(l.ref_shared_logf(-1.0)) and l.ref_shared_logf(1.0) == 0.0 and l.ref_shared_expf(0.0) == 1.0


def test_rolling_shutter_camera_vs_reference_glm(ref):
    """get_xform_given_rolling_shutter (common_device.cuh:401-408), the camera EVERY training ray of the reference is built from: position lerp +
    GLM quat_cast / slerp / normalize / mat3_cast.  tests/golden/rolling_shutter.npz holds the mount's own function (through the mount's GLM) on
    3000 seeded cases: identical start / end poses (the matrix -> quaternion -> matrix round trip alone moves the rotation by up to 6.6e-7),
    nearly identical ones (slerp's linear branch) and clearly different ones (acos / sin branch, libm on both sides).  Bit-identical."""
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "rolling_shutter.npz"))
    l = ref.lib()
    f32, vp = C.c_float, C.c_void_p
    l.ref_xform_given_rolling_shutter.argtypes = [vp, vp, vp, f32, f32, f32, vp]
    l.ref_xform_given_rolling_shutter.restype = None
    n = len(g["start"])
    got = np.zeros((n, 12), np.float32)
    s, e, rs, uvt = (np.ascontiguousarray(g[k]) for k in ("start", "end", "rolling_shutter", "uvt"))
    for k in range(n):
        l.ref_xform_given_rolling_shutter(s[k].ctypes.data_as(vp), e[k].ctypes.data_as(vp), rs[k].ctypes.data_as(vp), float(uvt[k, 0]), float(uvt[k, 1]),
                                          float(uvt[k, 2]), got[k].ctypes.data_as(vp))
    assert np.array_equal(got.view(np.uint32), g["xform"].view(np.uint32))
    assert np.abs(got[: n // 3, :9] - s[: n // 3, :9]).max() > 1e-7            # the round trip is not the identity: it is part of the ray


def test_rotmat_rotvec_vs_reference_glm(ref):
    """rotmat(vec3) / rotvec(mat3) (common_device.cuh:698-722), the two functions RotationAdamOptimizer::step composes a rotation step with
    (adam_optimizer.h:246): the oracle's restatement bit for bit against the reference's own code over 4000 angle-axis vectors from 1e-4 rad up,
    and through the optimiser's composition rotvec(rotmat(-step) * rotmat(variable)).  Golden file: tests/golden/make_golden_rotvec.py."""
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "rotvec.npz"))
    f = np.float32
    for k in range(len(g["vec"])):
        m = ref.rotmat(g["vec"][k])
        assert np.array_equal(m.T.reshape(-1), g["mat"][k]), k
        assert np.array_equal(ref.rotvec(g["mat"][k].reshape(3, 3).T), g["back"][k]), k
    assert np.array_equal(g["mat"][0], np.eye(3, dtype=np.float32).reshape(-1))            # the zero vector: exactly the identity
    for k in range(0, len(g["step"]), 7):
        A, B = ref.rotmat(-g["step"][k]), ref.rotmat(g["var"][k])
        prod = np.zeros((3, 3), np.float32)
        for c in range(3):
            for r in range(3):
                prod[r, c] = f(f(f(A[r, 0] * B[0, c]) + f(A[r, 1] * B[1, c])) + f(A[r, 2] * B[2, c]))
        assert np.array_equal(ref.rotvec(prod), g["composed"][k]), k
    # glm's axis() divides by sqrt(1 - w^2) in float32: at the size of one optimiser step (1e-3 rad) the round trip is only good to a few
    # percent, and below ~3e-4 rad w rounds to 1 and the axis degenerates to (0, 0, 1).  That is the reference's arithmetic; it is restated, not repaired.
    n = np.linalg.norm(g["vec"], axis=1)
    rel = np.linalg.norm(g["back"] - g["vec"], axis=1) / np.maximum(n, 1e-30)
    step_sized = (n > 1e-3) & (n < 3e-3)
    assert 1e-3 < rel[step_sized].max() < 0.5
    assert rel[(n > 0) & (n < 2e-4)].max() > 0.5
    assert rel[(n > 0.05) & (n < 0.9)].max() < 1e-3
