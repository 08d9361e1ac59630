"""Generates tests/golden/reference_helpers.npz from the reference's OWN helper code (oracle/_ref/libngp_refshim.so, built
by `make -C oracle ref` from /root/reference; see oracle/ref_shim/shim.cpp).  Run in the container that has the mount:

    make -C oracle ref && python tests/golden/make_golden.py

The .npz holds seeded inputs and the reference's outputs; tests/test_oracle_pins.py checks the oracle (and, through the
oracle-parity GPU tests, the CUDA kernels) against it on machines where /root/reference does not exist."""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
shim = C.CDLL(os.path.join(HERE, "..", "..", "oracle", "_ref", "libngp_refshim.so"))
f32, u32, i32, vp = C.c_float, C.c_uint32, C.c_int, C.c_void_p
for name, res, args in [("shim_ld_random_val", f32, [u32, u32]), ("shim_srgb_to_linear", f32, [f32]), ("shim_linear_to_srgb", f32, [f32]),
                        ("shim_to_stepping_space", f32, [f32, f32]), ("shim_from_stepping_space", f32, [f32, f32]), ("shim_advance_n_steps", f32, [f32, f32, f32]),
                        ("shim_calc_dt", f32, [f32, f32]), ("shim_advance_to_next_voxel", f32, [f32, f32, vp, vp, u32]), ("shim_mip_from_pos", u32, [vp, u32]),
                        ("shim_mip_from_dt", u32, [f32, vp, u32]), ("shim_cascaded_grid_idx_at", u32, [vp, u32]), ("shim_density_grid_occupied_at", i32, [vp, vp, u32]),
                        ("shim_warp_dt", f32, [f32]), ("shim_unwarp_dt", f32, [f32]), ("shim_network_to_rgb", f32, [f32, i32]),
                        ("shim_network_to_rgb_derivative", f32, [f32, i32]), ("shim_network_to_density", f32, [f32, i32]),
                        ("shim_network_to_density_derivative", f32, [f32, i32]), ("shim_aabb_contains", i32, [vp, vp, vp])]:
    getattr(shim, name).restype = res; getattr(shim, name).argtypes = args
P = lambda a: a.ctypes.data_as(vp)
rng = np.random.default_rng(20261104)
N = 4096
out = {}
# low-discrepancy sequence (random_val.cuh:162-291)
idx = rng.integers(0, 2 ** 32, N, dtype=np.uint64).astype(np.uint32); seeds = rng.integers(0, 2 ** 32, N, dtype=np.uint64).astype(np.uint32)
out["ld_index"], out["ld_seed"] = idx, seeds
out["ld_val"] = np.array([shim.shim_ld_random_val(int(a), int(b)) for a, b in zip(idx, seeds)], np.float32)
v2 = np.zeros((N, 2), np.float32)
for k in range(N):
    shim.shim_ld_random_val_2d(u32(int(idx[k])), u32(int(seeds[k])), P(v2[k]))
out["ld_val_2d"] = v2
# colour (common_device.cuh:75-121)
x = rng.uniform(0, 1.2, N).astype(np.float32); out["colour_x"] = x
out["srgb_to_linear"] = np.array([shim.shim_srgb_to_linear(float(a)) for a in x], np.float32)
out["linear_to_srgb"] = np.array([shim.shim_linear_to_srgb(float(a)) for a in x], np.float32)
# stepping (testbed_nerf.cu:124-184): cone angle 0 (aabb_scale 1) and 1/256 (aabb_scale > 1)
t = np.concatenate([rng.uniform(0, 4, N // 2), rng.uniform(0, 200, N // 2)]).astype(np.float32); out["step_t"] = t
for tag, cone in (("c0", 0.0), ("c256", 1.0 / 256.0)):
    out[f"to_stepping_{tag}"] = np.array([shim.shim_to_stepping_space(float(a), cone) for a in t], np.float32)
    out[f"from_stepping_{tag}"] = np.array([shim.shim_from_stepping_space(float(a), cone) for a in out[f"to_stepping_{tag}"]], np.float32)
    out[f"calc_dt_{tag}"] = np.array([shim.shim_calc_dt(float(a), cone) for a in t], np.float32)
    out[f"advance_half_{tag}"] = np.array([shim.shim_advance_n_steps(float(a), cone, 0.5) for a in t], np.float32)
# occupancy look-ups + voxel skip (testbed_nerf.cu:285-315,433-457,614-633)
pos = rng.uniform(-3.5, 4.5, (N, 3)).astype(np.float32); pos[: N // 2] = rng.uniform(0, 1, (N // 2, 3)).astype(np.float32)
d = rng.normal(0, 1, (N, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True); d = d.astype(np.float32)
bitfield = rng.integers(0, 256, 128 ** 3 * 8 // 8, dtype=np.uint16).astype(np.uint8)
out["occ_pos"], out["occ_dir"], out["bitfield_seed"] = pos, d, np.array([77], np.uint32)
bitfield = np.random.default_rng(77).integers(0, 256, 128 ** 3 * 8 // 8, dtype=np.uint16).astype(np.uint8)
tt = rng.uniform(0.1, 3, N).astype(np.float32); out["occ_t"] = tt
for mc in (0, 3, 7):
    out[f"mip_from_pos_{mc}"] = np.array([shim.shim_mip_from_pos(P(p), mc) for p in pos], np.uint32)
    out[f"mip_from_dt_{mc}"] = np.array([shim.shim_mip_from_dt(float(dt), P(p), mc) for p, dt in zip(pos, out["calc_dt_c256"])], np.uint32)
mips = out["mip_from_pos_7"]
out["grid_idx"] = np.array([shim.shim_cascaded_grid_idx_at(P(p), int(m)) for p, m in zip(pos, mips)], np.uint32)
out["occupied"] = np.array([shim.shim_density_grid_occupied_at(P(p), P(bitfield), int(m)) for p, m in zip(pos, mips)], np.int32)
for tag, cone in (("c0", 0.0), ("c256", 1.0 / 256.0)):
    out[f"next_voxel_{tag}"] = np.array([shim.shim_advance_to_next_voxel(float(a), cone, P(p), P(q), int(m)) for a, p, q, m in zip(tt, pos, d, mips)], np.float32)
# bounding box (bounding_box.cuh:163-223)
mn, mx = np.zeros(3, np.float32), np.ones(3, np.float32)
o = rng.uniform(-2, 3, (N, 3)).astype(np.float32); out["aabb_o"] = o
isect = np.zeros((N, 2), np.float32)
for k in range(N):
    shim.shim_aabb_ray_intersect(P(mn), P(mx), P(o[k]), P(d[k]), P(isect[k]))
out["aabb_isect"] = isect
out["aabb_contains"] = np.array([shim.shim_aabb_contains(P(mn), P(mx), P(p)) for p in pos], np.int32)
# dt warps, activations, losses (testbed_nerf.cu:186-283,317-431,1340-1355)
dtv = rng.uniform(0.001, 0.3, N).astype(np.float32); out["dt"] = dtv
out["warp_dt"] = np.array([shim.shim_warp_dt(float(a)) for a in dtv], np.float32)
out["unwarp_dt"] = np.array([shim.shim_unwarp_dt(float(a)) for a in out["warp_dt"]], np.float32)
v = rng.normal(0, 6, N).astype(np.float32); out["act_x"] = v
for act in range(4):
    for fn in ("network_to_rgb", "network_to_rgb_derivative", "network_to_density", "network_to_density_derivative"):
        out[f"{fn}_{act}"] = np.array([getattr(shim, "shim_" + fn)(float(a), act) for a in v], np.float32)
tg = rng.uniform(0, 1, (N, 3)).astype(np.float32); pr = (tg + rng.normal(0, 0.2, (N, 3))).astype(np.float32); out["loss_target"], out["loss_pred"] = tg, pr
for lt in range(7):
    lg = np.zeros((N, 6), np.float32)
    for k in range(N):
        shim.shim_loss_and_gradient(P(tg[k]), P(pr[k]), lt, P(lg[k]))
    out[f"loss_{lt}"] = lg
np.savez_compressed(os.path.join(HERE, "reference_helpers.npz"), **out)
print("wrote", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "reference_helpers.npz")), "bytes")
