import importlib, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
pyngp = importlib.import_module("instant-ngp_b200.pyngp"); scene = importlib.import_module("instant-ngp_b200.scene")
sys.path.insert(0, "tests")
import tests.test_render_frame_gpu as T
import tests.test_nerf_gpu as N
for name, cfg, nv, res in (("frame_cfg", T.CFG, 8, 64), ("small_cfg", N.SMALL_CFG, 8, 64), ("frame_cfg_12x96", T.CFG, 12, 96)):
    views = scene.make_views(nv, res, res)
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf); tb.set_nerf_dataset(views, aabb_scale=1); tb.reload_network_from_json(cfg)
    out = {}
    for steps in (150, 150, 300):
        for _ in range(steps): tb.train(1 << 14)
        ev = tb.evaluate_training_views(None, spp=8)
        out[tb.training_step] = (round(tb.loss, 5), round(ev["psnr"], 2), round(ev["ssim"], 3), tb.stats.rays_per_batch)
    print(name, out)
