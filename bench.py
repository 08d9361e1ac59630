#!/usr/bin/env python
"""bench.py — NeRF training throughput of the B200-native instant-ngp hot path (BASELINE.json metric:
"NeRF train samples/s & ms/step at 1/2/4/8 B200").

    python bench.py --gpus 1 --steps K --warmup W                 # own arm (CUDA, sm_100a)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W                 # data parallel, one rank per GPU, NCCL
    python bench.py --impl reference --gpus N --steps K --warmup W  # CPU arm: the oracle's restatement of the same step

A "step" is one Testbed::train(2^18) call (testbed.cu:4285-4370): [occupancy-grid refresh at its cadence] -> ray
sampler -> network inference on the un-compacted samples -> compositing loss + compaction -> forward/backward on the
compacted batch -> fused Adam(+EMA) -> counter read-back.  Workload: 100 posed 800x800 RGBA8 views of the procedural
analytic density field (north_star), HashGrid 16x2 T=2^19 + 1x64 density / 2x64 rgb FullyFusedMLP, Huber loss.
value = compacted samples trained per second, summed over ranks (dataset resident in HBM);
e2e   = the same metric through the pyngp-shaped API with a HOST->device upload of one training view per step
        (nerf.training.set_image, python_api.cu:49-67) and the per-step counter/loss read-back inside the timed region.
Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import importlib
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "nerf_train_samples_per_s"
UNIT = "samples/s"
N_VIEWS, RES = 100, 800
BATCH = 1 << 18
PRETRAIN = 1024         # untimed steps before warm-up: the occupancy grid has sparsified and rays_per_batch has settled (tools/phase_profile.py)


def network_config(log2_hashmap_size=19):
    return {
        "loss": {"otype": "Huber"},
        "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                      "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
        "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": log2_hashmap_size, "base_resolution": 16,
                     "per_level_scale": 1.3819},
        "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
        "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]},
        "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
    }


def workload_name(args):
    return (f"nerf_synthetic_{args.views}x{args.res}x{args.res}_hashgrid16x2_T{args.log2_hashmap_size}_mlp1x64+2x64_batch2^{int(math.log2(args.batch))}")


# ---------------------------------------------------------------------------------------------------------------
# clocks: sampled with nvidia-smi DURING the timed region (B200_PROFILING.md recipe)
# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, windows=None):
        """median SM clock over the samples that arrived inside the given host-time windows (the timed regions); the sampler runs
        from the untimed pre-training on, i.e. every sample is taken under the same training load"""
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        sm_in, sm_all, mx, reasons = [], [], 0, set()
        for ts, l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 7:
                continue
            try:
                v = float(f[0]); mx = max(mx, float(f[1]))
            except ValueError:
                continue
            sm_all.append(v)
            if windows and any(a <= ts <= b + 0.03 for a, b in windows):
                sm_in.append(v)
            for n, x in zip(names, f[3:7]):
                if x.lower().startswith("active"):
                    reasons.add(n)
        use = sorted(sm_in if sm_in else sm_all)
        return {"sm_mhz": use[len(use) // 2] if use else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm_in),
                "samples_under_load": len(sm_all)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "tflops_burst": float(d["bf16_tflops"]), "tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def profiled_traffic():
    """dram bytes per launch of the top kernels from the committed ncu --set full capture (profiles/traffic.json), else {}"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm (oracle): bounded sample of the same step
# ---------------------------------------------------------------------------------------------------------------
def cpu_arm(steps, warmup, batch, n_views, res, log2T):
    """Times `steps` Testbed::train steps of oracle/cpu_train.py (all host threads via OpenMP).
    Sample: batch `batch` on `n_views` views of res x res of the same analytic scene and the same network config;
    3 start-up steps, then the step counter is moved to the steady-state cadence (grid refresh every 16 steps over
    128^3/2 cells, testbed.cu:4321 / testbed_nerf.cu:4146-4150) so that the refresh is amortised as in a long run."""
    from oracle import cpu_train, ref_api as ref
    cores = os.cpu_count() or 1
    focal = 0.5 * res / math.tan(0.5 * 0.6911)
    views = ref.synth_views(n_views, res, res, 4.0 * 0.33, focal)
    tr = cpu_train.CpuNerfTrainer(network_config(log2T), views)
    for _ in range(3):
        tr.train(batch)
    tr.training_step = 4096 - warmup      # the timed region starts on a grid-refresh step and holds one refresh per 16 steps
    for _ in range(warmup):
        tr.train(batch)
    tr.timings.clear()
    n = 0
    t0 = time.perf_counter()
    for _ in range(steps):
        n += tr.train(batch)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{steps} steps of batch 2^{int(math.log2(batch))} ({n} samples, {dt:.1f} s) on {n_views} views {res}x{res}, "
                      f"OpenMP over {cores} threads, steady-state grid cadence; samples before compaction/step ~{tr.measured_before}",
            "ms_per_step": dt / steps * 1e3, "seconds": dt, "breakdown_s": {k: round(v, 3) for k, v in tr.timings.items()}}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 16)); warm = max(0, min(args.warmup, 2))
    r = cpu_arm(steps, warm, 1 << 12, 8, 200, args.log2_hashmap_size)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 storage / f32 accumulate",
            "data": "synthetic", "config": {"workload": workload_name(args), "note": "CPU restatement (oracle/) of the same training step; the reference has no CPU "
                                            "implementation and its CUDA build needs the absent tiny-cuda-nn submodule"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
# own arm
# ---------------------------------------------------------------------------------------------------------------
def algorithmic(kernel, cfg):
    """(bound, per-unit algorithmic bytes or flops, unit name) of a profiled kernel group."""
    enc = cfg["encoding"]
    L, F, D = enc["n_levels"], enc["n_features_per_level"], 3
    Cw = (L * F + 15) // 16 * 16
    dh, rh = cfg["network"]["n_hidden_layers"], cfg["rgb_network"]["n_hidden_layers"]
    f_density = 2 * (Cw * 64 + (dh - 1) * 64 * 64 + 64 * 16)
    f_rgb = 2 * (32 * 64 + (rh - 1) * 64 * 64 + 64 * 16)
    grid_bytes = 4 * D + L * (1 << D) * F * 2 + Cw * 2
    table = {
        "grid_forward": ("hbm", grid_bytes, "point"),
        "grid_backward": ("hbm", grid_bytes, "point"),
        "nerf_mlp_forward": ("tensor", f_density + f_rgb, "sample"),
        "nerf_mlp_density": ("tensor", f_density, "sample"),
        "nerf_mlp_backward": ("tensor", 3 * (f_density + f_rgb), "sample"),
        "adam": ("hbm", 30, "param"),
        "sampler": ("hbm", None, "ray"), "loss": ("hbm", None, "ray"), "fill_rollover": ("hbm", 36, "sample"),
    }
    return table.get(kernel)


def run_own(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the product path has no CPU fallback"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    _lib = importlib.import_module("instant-ngp_b200._lib")
    lib = _lib.lib()
    cfg = network_config(args.log2_hashmap_size)

    views = scene.make_views(args.views, args.res, args.res)            # pinned HOST buffers (like decoded images)
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(views, aabb_scale=1)
    tb.reload_network_from_json(cfg)
    if world > 1:
        tb.enable_data_parallel(native=os.environ.get('NGP_DP_NATIVE', '1') == '1')
    batch = args.batch

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clk = ClockSampler(local).__enter__()      # runs from the untimed pre-training to the end of the e2e pass: every sample under the same training load
    for _ in range(args.pretrain + args.warmup):
        tb.train(batch)

    def timed(steps, upload):
        """K steps; returns (ms max over ranks, samples summed over ranks, samples before compaction, launches)"""
        n = 0; nb = 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        l0 = lib.ngp_kernel_launch_count()
        e0.record()
        for i in range(steps):
            if upload:
                tb.set_training_image(i % len(views), views[i % len(views)]["pixels"])    # H2D from pinned host memory
            tb.train(batch)
            st = tb.stats                                                                  # host copy of the counters read back by train()
            n += st.measured_batch_size; nb += st.measured_batch_size_before_compaction
            _ = st.loss
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = lib.ngp_kernel_launch_count() - l0
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
            c = torch.tensor([n, nb], dtype=torch.float64, device="cuda"); dist.all_reduce(c); n, nb = int(c[0].item()), int(c[1].item())
        return ms, n, nb, launches

    ncu = os.environ.get("NGP_NCU") == "1"      # under `ncu --profile-from-start off`: capture only the timed steps
    if ncu:
        torch.cuda.profiler.start()
    t0 = time.time()
    ms, n_samples, n_before, launches = timed(args.steps, upload=False)
    t1 = time.time()
    if ncu:
        torch.cuda.profiler.stop()
    clk.__exit__()                              # (NVML polling during the upload pass was measured to slow the H2D copies: stop it first)
    clocks = clk.summary([(t0, t1)])
    time.sleep(0.25)                            # let the stopped NVML poller go away
    timed(max(args.warmup, 3), upload=True)     # untimed warm-up of the upload path (copy stream, events, first pinned-page touches)
    ms_e2e, n_e2e, _, _ = timed(args.steps, upload=True)
    view_bytes = int(views[0]["pixels"].numel())
    loss_steps = sum(1 for i in range(args.steps) if i % 16 == 0)
    d2h = 16 + 4.0 * loss_steps / args.steps

    # per-kernel CUDA-event timing on the launching stream over a third pass of K steps (events add ~2 records per kernel,
    # so this pass is not the one `value` comes from)
    prefetch_hits = int(tb.prefetch_hits)
    tb.prefetch_rays = False            # kernels timed one at a time: no ray prefetch running beside them on the second stream
    lib.ngp_profile_enable(1)
    buf = C.create_string_buffer(1 << 16)
    lib.ngp_profile_report(buf, len(buf), 1)
    timed(args.steps, upload=False)
    lib.ngp_profile_report(buf, len(buf), 1)
    lib.ngp_profile_enable(0)
    prof = json.loads(buf.value.decode())
    peaks = measured_peaks()
    traffic = profiled_traffic()
    kernels = {}
    nested = {"train_nerf_step", "density_grid_update"}      # enclosing scopes, not kernels
    for name, r in prof.items():
        if name in nested or r["count"] == 0:
            continue
        a = algorithmic(name, cfg)
        row = {"launches_per_step": r["count"] / args.steps, "ms_per_step": r["ms"] / args.steps, "avg_ms": r["ms"] / r["count"], "units_per_launch": r["units"] / r["count"]}
        if a and a[1]:
            per_launch = a[1] * r["units"] / r["count"]
            ach = per_launch / (r["ms"] / r["count"] * 1e-3)
            if a[0] == "hbm":
                row.update(bound="hbm", achieved=ach / 1e9, peak=peaks["hbm_gbs"], unit="GB/s", frac=ach / 1e9 / peaks["hbm_gbs"], per_unit=a[1])
            else:
                row.update(bound="tensor", achieved=ach / 1e12, peak=peaks["tflops_sustained"], unit="TFLOP/s", frac=ach / 1e12 / peaks["tflops_sustained"], per_unit=a[1])
            row["traffic"] = traffic.get(name)
        kernels[name] = row
    step_ms_prof = sum(k["ms_per_step"] for k in kernels.values())
    cand = [k for k in kernels if "achieved" in kernels[k]]
    top = max(cand, key=lambda k: kernels[k]["ms_per_step"])
    roof = {k: kernels[top][k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic")}
    roof.update(kernel=top, share_of_step=kernels[top]["ms_per_step"] / step_ms_prof, avg_ms=kernels[top]["avg_ms"],
                peak_source=peaks["source"] + (", sustained bf16" if roof["bound"] == "tensor" else ", copy bandwidth"))

    line = {"metric": METRIC, "value": n_samples / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 (fp32 accumulate, fp32 master weights)",
            "data": "synthetic",
            "config": {"workload": workload_name(args), "global_batch": batch * world, "per_gpu_batch": batch, "parallelism": f"dp{world}" if world > 1 else "single",
                       "pretrain_steps_untimed": args.pretrain, "samples_before_compaction_per_step": n_before / args.steps / world,
                       "rays_per_batch": tb.stats.rays_per_batch, "n_params": int(tb.n_params()), "steps_with_prefetched_rays": prefetch_hits,
                       "l2_policy": "inputs larger than L2: 256 MB image set + 24 MB table + >100 MB sample buffers touched every step"},
            "e2e": {"value": n_e2e / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": view_bytes, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps,
                    "api": "pyngp.Testbed.set_training_image + train + loss"},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "kernels": kernels}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_arm(8, 1, 1 << 12, 8, 200, args.log2_hashmap_size)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pretrain", type=int, default=PRETRAIN)
    ap.add_argument("--views", type=int, default=N_VIEWS)
    ap.add_argument("--res", type=int, default=RES)
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--log2-hashmap-size", type=int, default=19)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
