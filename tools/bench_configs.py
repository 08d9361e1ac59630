"""Secondary BASELINE configs on one GPU (tools): config 2 (neural image, 2-D HashGrid T=2^24 + 2x64 MLP, 2^18 pixels/step on a
synthetic 1024x1024 fp32 image — data/image/albert.exr is not on the GPU box) and config 5's single-GPU leg (NeRF, T=2^24).
CUDA events over K steps after warm-up; prints one JSON object."""
import importlib, json, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
pyngp = importlib.import_module("instant-ngp_b200.pyngp"); scene = importlib.import_module("instant-ngp_b200.scene")
res = {}

def timed(fn, steps, warm):
    for _ in range(warm): fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(steps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / steps

# ---- config 2: neural image ----
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
    N = 1 << 18
    ms = timed(lambda: tb.train(N), 200, 800)
    mse = tb.compute_image_mse()
    res[f"config2_image_b{b}"] = {"ms_per_step": ms, "Mpixels_s": N / ms / 1e3, "n_params": int(tb.n_params()), "psnr_after_1000_steps": -10 * math.log10(max(mse, 1e-12))}
    del tb

# ---- config 5 (one GPU): NeRF with a 2^24-entry table ----
views = scene.make_views(100, 800, 800)
tb = pyngp.Testbed(pyngp.TestbedMode.Nerf); tb.set_nerf_dataset(views, aabb_scale=1); tb.reload_network_from_json(bench.network_config(24))
N = 1 << 18
n = [0]
def step():
    tb.train(N); n[0] += tb.stats.measured_batch_size
for _ in range(600): tb.train(N)
n[0] = 0
ms = timed(step, 64, 0)
res["config5_nerf_T24_1gpu"] = {"ms_per_step": ms, "samples_s": n[0] / 64 / ms * 1e3, "n_params": int(tb.n_params())}
print(json.dumps(res, indent=1))
json.dump(res, open("gpurun_out/configs.json", "w"), indent=1)
