import importlib, sys, os
sys.path.insert(0, '/root/repo')
import numpy as np, torch, bench
pyngp = importlib.import_module("instant-ngp_b200.pyngp"); scene = importlib.import_module("instant-ngp_b200.scene")
views = scene.make_views(100, 800, 800)
tb = pyngp.Testbed(pyngp.TestbedMode.Nerf); tb.set_nerf_dataset(views, aabb_scale=1); tb.reload_network_from_json(bench.network_config(19))
tb.prefetch_rays = False
for _ in range(1100): tb.train(1 << 18)
c = tb.debug_tensor("counters", torch.int32, 8).cpu().numpy()
ns = tb.debug_tensor("numsteps_all", torch.int32, tb.stats.rays_per_batch).cpu().numpy()
print("counters", c, "rays_per_batch", tb.stats.rays_per_batch, "rays with samples", (ns > 0).sum(), "max", ns.max(), "p99", np.percentile(ns[ns>0], 99))
