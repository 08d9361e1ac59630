#!/usr/bin/env python
"""GPU timeline of a few steady-state training steps with the ray prefetch ON (tools, not product): start and duration of every kernel
group on both streams, from the library's profiler events.  Shows where the training stream idles.  usage: timeline.py [n_steps]"""
import ctypes as C, importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
lib = importlib.import_module("instant-ngp_b200._lib").lib()
class A: views, res, log2_hashmap_size, batch = 100, 800, 19, 1 << 18
tb, views = bench.make_testbed(A)
for _ in range(1100):
    tb.train(A.batch)
torch.cuda.synchronize()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
buf = C.create_string_buffer(1 << 20)
lib.ngp_profile_enable(1); lib.ngp_profile_report(buf, len(buf), 1)
for _ in range(n):
    tb.train(A.batch)
torch.cuda.synchronize()
lib.ngp_profile_timeline(buf, len(buf)); lib.ngp_profile_enable(0)
tl = json.loads(buf.value.decode())
steps = [r for r in tl if r[0] == "train_nerf_step"]
print(f"{n} steps, {len(tl)} records; step starts (us): {[round(s[1]) for s in steps]}")
t_end_prev = None
for name, t0, d in sorted(tl, key=lambda r: r[1]):
    if name in ("train_nerf_step", "density_grid_update"):
        print(f"{t0:10.1f}  ---- {name} (host scope {d:.1f} us)")
        continue
    stream = "aux " if name == "sampler" else "main"
    gap = "" if (t_end_prev is None or stream != "main") else f"   idle before: {t0 - t_end_prev:6.1f}"
    print(f"{t0:10.1f}  {stream} {name:18s} {d:7.1f} us{gap}")
    if stream == "main":
        t_end_prev = t0 + d
