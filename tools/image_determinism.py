import importlib, json, math, os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
pyngp = importlib.import_module("instant-ngp_b200.pyngp")
W = H = 1024
y, x = np.mgrid[0:H, 0:W].astype(np.float32)
img = np.zeros((H, W, 4), np.float32)
img[..., 0] = 0.5 + 0.5 * np.sin(x / 9.0) * np.cos(y / 13.0); img[..., 1] = ((x // 16 + y // 16) % 2) * 0.8 + 0.1
img[..., 2] = np.clip((x + y) / (W + H), 0, 1); img[..., 3] = 1.0
for b in (2.0, 1.3819):
    cfg = {"loss": {"otype": "L2"},
           "optimizer": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                         "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}},
           "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24, "base_resolution": 16, "per_level_scale": b},
           "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}}
    tb = pyngp.Testbed(pyngp.TestbedMode.Image); tb.set_image(img); tb.reload_network_from_json(cfg)
    out = []
    for k in range(1, 1201):
        tb.train(1 << 18)
        if os.environ.get("SYNC") == "1": torch.cuda.synchronize()
        if k in (200, 600, 1000, 1200):
            out.append(round(-10 * math.log10(max(tb.compute_image_mse(), 1e-12)), 2))
    print("agg", os.environ.get("NGP_AGG_MAX_RUNS"), "sync", os.environ.get("SYNC"), "b", b, "psnr@200/600/1000/1200", out, flush=True)
