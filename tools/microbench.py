"""Developer micro-benchmark of the operator-level kernels (CUDA events, L2 flushed between iterations)."""
import importlib, json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
importlib.import_module("instant-ngp_b200"); ops = importlib.import_module("instant-ngp_b200.ops")

def timeit(fn, iters=20, warmup=3, flush=None):
    for _ in range(warmup): fn()
    ts = []
    for _ in range(iters):
        if flush is not None: flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(np.min(ts))

def main():
    N = 1 << 18
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    res = {}
    for name, b, T in [("16x2_b1.38_T19", 1.3819, 19), ("16x2_b2_T19", 2.0, 19), ("16x2_b1.38_T24", 1.3819, 24)]:
        enc = ops.Encoding({"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": T, "base_resolution": 16, "per_level_scale": b}, 3)
        table = (torch.rand(enc.n_params(), device="cuda") * 2e-4 - 1e-4).half()
        x = torch.rand(N, 3, device="cuda")
        dy = (torch.randn(N, 32, device="cuda") / 128).half()
        grad = torch.zeros(enc.n_params(), dtype=torch.float16, device="cuda")
        y = torch.empty(N * 32, dtype=torch.float16, device="cuda")
        import ctypes as C
        from importlib import import_module
        L = import_module("instant-ngp_b200._lib").lib()
        s = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
        fwd = lambda: L.ngp_encoding_forward(enc._h, s(), C.c_void_p(table.data_ptr()), C.c_void_p(x.data_ptr()), 3, N, C.c_void_p(y.data_ptr()), 1)
        bwd = lambda: L.ngp_encoding_backward(enc._h, s(), C.c_void_p(x.data_ptr()), 3, N, C.c_void_p(dy.data_ptr()), 0, C.c_void_p(grad.data_ptr()))
        for tag, fn in [("fwd", fwd), ("bwd", bwd)]:
            cold = timeit(fn, flush=flush); warm = timeit(fn)
            res[f"grid_{name}_{tag}"] = {"ms_cold_med": cold[0], "ms_warm_med": warm[0], "ms_warm_min": warm[1],
                                        "Gpoints_s_warm": N / warm[0] / 1e6, "GBs_alg_warm": N * 588 / warm[0] / 1e6}
    # denominator for "fraction of L2 gather bandwidth" (SURVEY 8d): random 4-byte gathers over a 24.4 MB (L2-resident) table,
    # torch.take as a neutral gather kernel; 1 sector (32 B) per gather
    tab = torch.randint(0, 1 << 30, (6098120,), dtype=torch.int32, device="cuda")
    idx = torch.randint(0, tab.numel(), (1 << 26,), dtype=torch.int64, device="cuda")
    t = timeit(lambda: torch.take(tab, idx), iters=10)
    res["l2_random_gather_4B"] = {"ms_med": t[0], "Ggathers_s": idx.numel() / t[0] / 1e6, "sector_GBs": idx.numel() * 32 / t[0] / 1e6, "table_MB": tab.numel() * 4 / 1e6}
    for k in list(res):
        if k.startswith("grid_") and "T19" in k:
            res[k]["Ggathers_s_warm"] = N * 128 / res[k]["ms_warm_med"] / 1e6
            res[k]["frac_of_l2_gather_rate"] = res[k]["Ggathers_s_warm"] / res["l2_random_gather_4B"]["Ggathers_s"]
    cfg = {"encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819},
           "network": {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": 1}, "rgb_network": {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": 2}}
    net = ops.NerfNetwork(cfg)
    params = net.initialize_params().half()
    coords = torch.rand(N, 7, device="cuda")
    dL = (torch.randn(N, 4, device="cuda") / 1024).half()
    grads = torch.empty(net.n_params(), dtype=torch.float16, device="cuda")
    t = timeit(lambda: net.inference(params, coords)); res["nerf_inference"] = {"ms_med": t[0], "ms_min": t[1], "Msamples_s": N / t[0] / 1e3}
    t = timeit(lambda: net.forward_backward(params, coords, dL, grads)); res["nerf_fwd_bwd"] = {"ms_med": t[0], "ms_min": t[1], "Msamples_s": N / t[0] / 1e3}
    print(json.dumps(res, indent=1))
    json.dump(res, open("gpurun_out/microbench.json", "w"), indent=1)

if __name__ == "__main__":
    main()
