#!/usr/bin/env python
"""Times the fused NeRF MLP kernels alone (tools, not product): nerf inference (hash grid + MLP forward) and forward_backward on random
inputs at the headline shape, per-kernel CUDA events of the library's profiler, mean of REPS.  Environment switches (NGP_MLP_TS, NGP_MLP_PROBE ...)
are read by the library at first use, so A/B = one process per setting.  Prints one JSON line."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ctypes as C
ops = importlib.import_module("instant-ngp_b200.ops")
lib = importlib.import_module("instant-ngp_b200._lib").lib()
scene = importlib.import_module("instant-ngp_b200.scene")
n_fwd, n_bwd, reps = 465 * 1024, 1 << 18, int(os.environ.get("REPS", "30"))
net = ops.NerfNetwork(scene.NERF_CONFIG_16x2)
g = torch.Generator(device="cuda").manual_seed(1)
params = (torch.rand(net.n_params(), device="cuda", generator=g) - 0.5).half()
coords = torch.rand(n_fwd, 7, device="cuda", generator=g)
dL = (torch.randn(n_bwd, 4, device="cuda", generator=g) / 64).half()
grads = torch.empty(net.n_params(), dtype=torch.float16, device="cuda")
buf = C.create_string_buffer(1 << 16)
def prof(fn, n):
    for _ in range(3): fn()
    lib.ngp_profile_enable(1); lib.ngp_profile_report(buf, len(buf), 1)
    for _ in range(n): fn()
    torch.cuda.synchronize()
    lib.ngp_profile_report(buf, len(buf), 1); lib.ngp_profile_enable(0)
    return {k: round(v["ms"] / v["count"], 5) for k, v in json.loads(buf.value.decode()).items() if v["count"]}
out = {"env": {k: v for k, v in os.environ.items() if k.startswith("NGP_")}}
out["inference_465k"] = prof(lambda: net.inference(params, coords), reps)
out["forward_backward_262k"] = prof(lambda: net.forward_backward(params, coords[:n_bwd], dL, grads), reps)
f = out["inference_465k"].get("nerf_mlp_forward"); b = out["forward_backward_262k"].get("nerf_mlp_backward")
out["tflops"] = {"fwd": round(20480 * n_fwd / (f * 1e-3) / 1e12, 1) if f else None, "bwd": round(61440 * n_bwd / (b * 1e-3) / 1e12, 1) if b else None}
print(json.dumps(out))
