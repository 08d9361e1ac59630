"""BASELINE config 3: NeRF fox (data/nerf/fox of the reference mount, 50 of 67 frames present, aabb_scale 8 -> 4 cascades, cone angle
1/256, OpenCV lens), 2^18 samples per step, trained to 35 000 steps (scripts/run.py:177-178), PSNR on the training views with the
run.py:208-266 protocol (black background, pixel centres, render_min_transmittance 1e-4, sRGB, clipped).

The scene is DATA of the reference, staged by hand into baseline/_data/fox (git-ignored; `cp -r /root/reference/data/nerf/fox
baseline/_data/fox` where the mount exists).  Prints one JSON line; {"unavailable": ...} when the data is not staged."""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

OPT = {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
       "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}}
MLPS = {"network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
        "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
        "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]}}
CONFIGS = {
    # the north-star network: 16 levels x 2 features, T = 2^19, growth to 2048 * aabb_scale as the reference derives it (testbed.cu:3985-3995)
    "hashgrid16x2": {"loss": {"otype": "Huber"}, "optimizer": OPT, **MLPS,
                     "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16,
                                  "per_level_scale": float(np.exp(np.log(2048 * 8 / 16) / 15))}},
    # the fork's configs/nerf/base.json: 4 levels x 4 features, per_level_scale left to its default
    "fork_base_4x4": {"loss": {"otype": "Huber"}, "optimizer": OPT, **MLPS,
                      "encoding": {"otype": "HashGrid", "n_levels": 4, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16}},
}


def psnr_on_training_views(tb, views, spp):
    return tb.evaluate_training_views(views, spp)["psnr"]      # scripts/run.py:208-266 protocol, see pyngp.Testbed.evaluate_training_views


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=35000)
    ap.add_argument("--config", default="hashgrid16x2", choices=sorted(CONFIGS))
    ap.add_argument("--spp", type=int, default=8)
    ap.add_argument("--eval-views", type=int, default=10)
    ap.add_argument("--profile", type=int, default=0, help="after training: per-kernel CUDA-event times over this many steps (prefetch off)")
    ap.add_argument("--data", default=os.path.join(ROOT, "baseline", "_data", "fox"))
    args = ap.parse_args()
    if not os.path.exists(os.path.join(args.data, "transforms.json")):
        print(json.dumps({"config": "nerf_fox", "unavailable": f"{args.data} is not staged (reference data, git-ignored)"}))
        return
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    t0 = time.time()
    tb.load_training_data(os.path.join(args.data, "transforms.json"))
    load_s = time.time() - t0
    tb.reload_network_from_json(CONFIGS[args.config])
    tb.sampler_lanes_per_ray = int(os.environ.get("NGP_LANES", "0"))
    if os.environ.get("NGP_WALK_AHEAD") == "0":
        tb.walk_ahead = False
    n_views = tb.n_training_views()
    views = list(range(0, n_views, max(n_views // args.eval_views, 1)))[:args.eval_views]
    checkpoints = [c for c in (1000, 5000, 35000) if c <= args.steps] or [args.steps]
    out = {"config": "nerf_fox", "network": args.config, "n_images": n_views, "batch": 1 << 18, "load_s": round(load_s, 2), "n_params": tb.n_params(),
           "psnr_protocol": f"{len(views)} training views, spp {args.spp}, black bg, min transmittance 1e-4, sRGB clipped", "checkpoints": {}}
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    step, samples_total, ms_total = 0, 0, 0.0
    for c in checkpoints:
        n = c - step
        samples = 0
        torch.cuda.synchronize(); ev0.record()
        for _ in range(n):
            tb.train(1 << 18)
            samples += tb.stats.measured_batch_size
        ev1.record(); torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        step = c; samples_total += samples; ms_total += ms
        st = tb.stats
        out["checkpoints"][str(c)] = {"ms_per_step": round(ms / n, 4), "samples_per_s": round(samples / ms * 1e3), "rays_per_batch": st.rays_per_batch,
                                      "samples_before_compaction": st.measured_batch_size_before_compaction, "loss": st.loss,
                                      "psnr_db": round(psnr_on_training_views(tb, views, args.spp), 3), "train_s_cumulative": round(ms_total / 1e3, 2)}
    if os.environ.get("NGP_NCU") == "1":      # capture window for `ncu --profile-from-start off`
        tb.prefetch_rays = False
        torch.cuda.synchronize(); torch.cuda.profiler.start()
        for _ in range(3):
            tb.train(1 << 18)
        torch.cuda.synchronize(); torch.cuda.profiler.stop()
    if args.profile:
        import ctypes as C
        lib = importlib.import_module("instant-ngp_b200._lib").lib()
        buf = C.create_string_buffer(1 << 16)
        for prefetch in (True, False):           # clean wall time first (no per-kernel events)
            tb.prefetch_rays = prefetch
            for _ in range(16 - tb.training_step % 16 + 2):
                tb.train(1 << 18)
            torch.cuda.synchronize(); ev0.record()
            for _ in range(64):
                tb.train(1 << 18)
            ev1.record(); torch.cuda.synchronize()
            out["ms_per_step_prefetch_" + ("on" if prefetch else "off")] = round(ev0.elapsed_time(ev1) / 64, 4)
        out["prefetch_hits"] = tb.prefetch_hits
        out["walk_ahead"] = {"enabled": bool(tb.walk_ahead), "hits": tb.walk_ahead_hits, "misses": tb.walk_ahead_misses}
        for prefetch in (True, False):
            tb.prefetch_rays = prefetch
            for _ in range(4):
                tb.train(1 << 18)
            lib.ngp_profile_enable(1); lib.ngp_profile_report(buf, len(buf), 1)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for _ in range(args.profile):
                tb.train(1 << 18)
            torch.cuda.synchronize(); wall = (time.perf_counter() - t0) / args.profile * 1e3
            lib.ngp_profile_report(buf, len(buf), 1); lib.ngp_profile_enable(0)
            prof = json.loads(buf.value.decode())
            out["profile_prefetch_" + ("on" if prefetch else "off")] = {"ms_per_step_wall": round(wall, 4),
                "kernels_ms_per_step": {k: round(v["ms"] / args.profile, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}}
    out["ms_per_step"] = round(ms_total / step, 4); out["samples_per_s"] = round(samples_total / ms_total * 1e3)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
