"""Per-kernel CUDA-event profile of the neural-image step (BASELINE config 2 shape) — tools, not product."""
import ctypes as C, importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
pyngp = importlib.import_module("instant-ngp_b200.pyngp"); lib = importlib.import_module("instant-ngp_b200._lib").lib()
W = H = 1024
y, x = np.mgrid[0:H, 0:W].astype(np.float32)
img = np.zeros((H, W, 4), np.float32)
img[..., 0] = 0.5 + 0.5 * np.sin(x / 9.0) * np.cos(y / 13.0); img[..., 1] = ((x // 16 + y // 16) % 2) * 0.8 + 0.1
img[..., 2] = np.clip((x + y) / (W + H), 0, 1); img[..., 3] = 1.0
b = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
cfg = {"loss": {"otype": "L2"},
       "optimizer": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                     "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}},
       "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24, "base_resolution": 16, "per_level_scale": b},
       "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}}
tb = pyngp.Testbed(pyngp.TestbedMode.Image); tb.set_image(img); tb.reload_network_from_json(cfg)
N = 1 << 18
for _ in range(200): tb.train(N)
buf = C.create_string_buffer(1 << 16)
lib.ngp_profile_enable(1); lib.ngp_profile_report(buf, len(buf), 1)
a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); a.record()
for _ in range(32): tb.train(N)
e.record(); torch.cuda.synchronize()
lib.ngp_profile_report(buf, len(buf), 1); lib.ngp_profile_enable(0)
prof = json.loads(buf.value.decode())
print(json.dumps({"b": b, "n_params": tb.n_params(), "ms_per_step_wall": a.elapsed_time(e) / 32, "kernels_ms_per_step": {k: round(v["ms"] / 32, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}}))
