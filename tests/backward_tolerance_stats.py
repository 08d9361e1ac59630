"""Margin of the hash-grid backward tolerance over repeated runs (fp16 atomics in run-dependent order) - a helper of tests/test_ops_gpu.py,
run by hand on a GPU box: python tests/backward_tolerance_stats.py"""
import sys, importlib
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from oracle import ref_api as ref
importlib.import_module("instant-ngp_b200"); ops = importlib.import_module("instant-ngp_b200.ops")
import tests.test_ops_gpu as T
for ci, case in enumerate(T.GRID_CASES):
    enc, cfg = T._grid(ops, ref, case)
    D, L, F = case[0], case[1], case[2]
    n = 16384
    rng = np.random.default_rng(11)
    x = ref.pcg32_floats(1338, 0, n * D).reshape(n, D)
    C = enc.padded_output_width()
    dy = np.zeros((n, C), np.float16)
    dy[:, :L * F] = (rng.normal(0, 1, (n, L * F)) / 128).astype(np.float16)
    dy[100:200] = 0
    gr = ref.grid_backward(cfg, x, dy, enc.n_params())
    gabs = ref.grid_backward(cfg, x, np.abs(dy), enc.n_params())
    tol = 2.0 ** -9 * gabs + 1e-6
    xs, dys = T._cuda(x), T._cuda(dy)
    worst = []
    for rep in range(40):
        g = enc.backward(xs, dys).float().cpu().numpy().astype(np.float64)
        bad = np.abs(g - gr) > tol
        ratio = (np.abs(g - gr) / tol).max()
        worst.append((int(bad.sum()), round(float(ratio), 2)))
    print("case", ci, case, "n_params", enc.n_params(), "allowed bad", 1e-4 * enc.n_params(), "max bad", max(w[0] for w in worst), "max ratio", max(w[1] for w in worst), flush=True)
