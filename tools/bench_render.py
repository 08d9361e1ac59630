"""Inference render timing (BASELINE config 5's render leg, one GPU's share and the full frame): the trained synthetic scene at
1920x1080 and 3840x2160, full frame and one of 8 horizontal bands, with per-kernel profile — tools, not product."""
import ctypes as C, importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
pyngp = importlib.import_module("instant-ngp_b200.pyngp"); scene = importlib.import_module("instant-ngp_b200.scene")
lib = importlib.import_module("instant-ngp_b200._lib").lib()
log2_T = int(sys.argv[1]) if len(sys.argv) > 1 else 19
views = scene.make_views(100, 800, 800)
tb = pyngp.Testbed(pyngp.TestbedMode.Nerf); tb.set_nerf_dataset(views, aabb_scale=1); tb.reload_network_from_json(bench.network_config(log2_T))
for _ in range(2000): tb.train(1 << 18)
torch.cuda.synchronize()
res = {"log2_hashmap_size": log2_T, "trained_steps": 2000}
buf = C.create_string_buffer(1 << 16)
for (w, h) in ((1920, 1080), (3840, 2160)):
    f = scene.focal_px(h) * 1.0
    out = torch.empty(h, w, 4, dtype=torch.float32, device="cuda")
    for name, (y0, y1) in (("full", (0, h)), ("band_1_of_8", (3 * h // 8, 4 * h // 8))):
        o = out[y0:y1]
        for _ in range(2): tb.render(w, h, 1, camera_matrix=views[7]["xform"], focal_length=(f, f), y0=y0, y1=y1, out=o)
        lib.ngp_profile_enable(1); lib.ngp_profile_report(buf, len(buf), 1)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        n = 5
        for _ in range(n): tb.render(w, h, 1, camera_matrix=views[7]["xform"], focal_length=(f, f), y0=y0, y1=y1, out=o)
        torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / n * 1e3
        lib.ngp_profile_report(buf, len(buf), 1); lib.ngp_profile_enable(0)
        prof = json.loads(buf.value.decode())
        res[f"{w}x{h}_{name}"] = {"ms_per_frame": round(ms, 3), "Mpixels_s": round(w * (y1 - y0) / ms / 1e3, 1), "coverage": float((o[..., 3] > 0.5).float().mean()),
                                  "kernels_ms": {k: round(v["ms"] / n, 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])[:8]},
                                  "launches": {k: v["count"] // n for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])[:8]}}
print(json.dumps(res))
