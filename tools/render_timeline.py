#!/usr/bin/env python
"""Per-iteration timeline of one 3840x2160 frame through the inference tracer (tools, not product): for every loop iteration the launch
bound and the duration of each kernel group, from the library's profiler events.  usage: render_timeline.py [log2_T] [W] [H]"""
import ctypes as C, importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
pyngp = importlib.import_module("instant-ngp_b200.pyngp"); scene = importlib.import_module("instant-ngp_b200.scene")
lib = importlib.import_module("instant-ngp_b200._lib").lib()
log2_T = int(sys.argv[1]) if len(sys.argv) > 1 else 19
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840; H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
views = scene.make_views(100, 800, 800)
tb = pyngp.Testbed(pyngp.TestbedMode.Nerf); tb.set_nerf_dataset(views, aabb_scale=1); tb.reload_network_from_json(bench.network_config(log2_T))
for _ in range(2000): tb.train(1 << 18)
f = scene.focal_px(H)
out = torch.empty(H, W, 4, dtype=torch.float32, device="cuda")
for _ in range(2): tb.render(W, H, 1, camera_matrix=views[7]["xform"], focal_length=(f, f), out=out)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(5): tb.render(W, H, 1, camera_matrix=views[7]["xform"], focal_length=(f, f), out=out)
e1.record(); torch.cuda.synchronize()
print(f"{W}x{H}: {e0.elapsed_time(e1) / 5:.3f} ms per frame without the profiler")
buf = C.create_string_buffer(1 << 22)
lib.ngp_profile_enable(1); lib.ngp_profile_report(buf, len(buf), 1)
tb.render(W, H, 1, camera_matrix=views[7]["xform"], focal_length=(f, f), out=out)
torch.cuda.synchronize()
lib.ngp_profile_timeline(buf, len(buf)); lib.ngp_profile_enable(0)
tl = sorted(json.loads(buf.value.decode()), key=lambda r: r[1])
t0 = tl[0][1]
it, row = 0, {}
print("iter   start_us  " + "  ".join(f"{k:>12s}" for k in ("compact", "next_inputs", "grid_forward", "mlp_forward", "composite")))
def flush():
    if row: print(f"{it:4d} {row.get('t', 0) - t0:10.1f}  " + "  ".join(f"{row.get(k, 0):12.1f}" for k in ("render_compact", "render_next_inputs", "grid_forward", "nerf_mlp_forward", "render_composite")))
for rec in tl:
    name, start, dur = rec[0], rec[1], rec[2]
    if name == "render_compact":
        flush(); it += 1; row = {"t": start}
    if name == "render_init_rays": print(f"init_rays {dur:.1f} us"); continue
    row[name] = row.get(name, 0) + dur
flush()
print(f"total span {tl[-1][1] + tl[-1][2] - t0:.1f} us over {it} iterations")
