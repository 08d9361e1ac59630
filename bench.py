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


def profiled_instructions():
    """warp instructions per unit of the issue-bound kernel groups, from the committed ncu capture (profiles/instructions.json), else {}"""
    p = os.path.join(ROOT, "profiles", "instructions.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


def profiled_traffic():
    """dram bytes per launch of the top kernels from the committed ncu --set full capture (profiles/traffic.json), else {}"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm (oracle): bounded sample of the same step, from the same trained state
# ---------------------------------------------------------------------------------------------------------------
def make_testbed(args, views=None):
    """the CUDA Testbed on the workload of `args` (dataset resident in HBM, network initialised)"""
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    if views is None:
        views = scene.make_views(args.views, args.res, args.res)            # pinned HOST buffers (like decoded images)
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(views, aabb_scale=1)
    tb.reload_network_from_json(network_config(args.log2_hashmap_size))
    return tb, views


def trained_state(tb):
    """host copy of everything a training step reads: parameters, optimizer moments are NOT needed for timing but the
    occupancy grid, bitfield, counters and rays_per_batch are — oracle/cpu_train.CpuNerfTrainer.adopt_state() continues from it"""
    import torch
    torch.cuda.synchronize()
    st = tb.stats
    return {"master": tb.param_buffer(0).cpu().numpy().copy(), "params": tb.param_buffer(1).cpu().numpy().copy(),
            "ema_h": tb.param_buffer(2).cpu().numpy().copy(),
            "density_grid": tb.density_grid(0).cpu().numpy().copy(), "bitfield": tb.density_grid(1).cpu().numpy().copy(),
            "mean_density": float(tb.debug_tensor("mean_density", torch.float32, 1).cpu().numpy()[0]),
            "training_step": int(st.training_step), "rays_per_batch": int(st.rays_per_batch),
            "measured_before": int(st.measured_batch_size_before_compaction), "measured": int(st.measured_batch_size)}


def cpu_arm(args, views, state, steps, warmup, batch, pretrained_how):
    """Times `steps` Testbed::train(batch) steps of oracle/cpu_train.py on all host threads (OpenMP), on the SAME views, network
    and — through `state` — the same trained occupancy grid / parameters / rays_per_batch as the GPU arm.  The occupancy-grid
    refresh (every 16th step at steady state, testbed.cu:4321) is timed once on its own and charged at 1/16 per step, so a
    sample of a few steps carries the cadence of a long run."""
    from oracle import cpu_train, ref_api as ref
    cores = os.cpu_count() or 1
    ref.lib().ref_set_num_threads(cores)                  # torchrun exports OMP_NUM_THREADS=1: set the count explicitly
    threads = int(ref.lib().ref_get_max_threads())
    tr = cpu_train.CpuNerfTrainer(network_config(args.log2_hashmap_size), views)
    if state is not None:
        tr.adopt_state(state)
    if state is not None:
        # move to the first step after a refresh at the steady-state cadence (16), keeping the adopted grid
        tr.training_step = max(tr.training_step, 4096) // 16 * 16 + 1
        tr.rng_skip = 1 + tr.training_step * (1 << 32)
    for _ in range(warmup if state is not None else max(warmup, 1)):   # (no adopted state: the first step builds the occupancy grid)
        tr.train(batch)
    if state is not None:
        tr.training_step = tr.training_step // 16 * 16 + 1     # timed steps: none of them refreshes (steps <= 15)
    tr.timings.clear()
    n = 0; before = 0
    t0 = time.perf_counter()
    for _ in range(steps):
        n += tr.train(batch); before += tr.measured_before
    dt = time.perf_counter() - t0
    t1 = time.perf_counter()
    tr.update_density_grid(cpu_train.NC // 4, cpu_train.NC // 4)       # one steady-state refresh (128^3/2 cells, :4146-4150)
    refresh = time.perf_counter() - t1
    sec_per_step = dt / steps + refresh / 16.0
    return {"value": (n / steps) / sec_per_step, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{steps} timed + {warmup} warm-up steps of batch 2^{int(math.log2(batch))} ({n} samples in {dt:.1f} s) + one occupancy-grid refresh "
                      f"({refresh:.2f} s, charged 1/16 per step) on {len(views)} views {args.res}x{args.res}, T=2^{args.log2_hashmap_size}; OpenMP threads {threads} of "
                      f"{cores} cores; state: {pretrained_how}; samples before compaction / trained sample {before / max(n, 1):.2f}",
            "ms_per_step": sec_per_step * 1e3, "seconds": dt + refresh, "threads": threads,
            "before_over_compacted": before / max(n, 1), "breakdown_s": {k: round(v, 3) for k, v in tr.timings.items()}}


def run_reference(args):
    """--impl reference: the reference has no CPU implementation and its CUDA build needs the absent tiny-cuda-nn submodule, so
    this arm times the oracle's restatement of the same training step on the host cores, on the SAME workload as the own arm
    (same views, network, batch) and from the same kind of trained state: when a GPU is present the state comes from `pretrain`
    untimed steps of the CUDA Testbed (data synthesis and pre-training only; nothing of the product runs inside the timed
    region), otherwise from a bounded number of CPU steps (said so in `config.workload`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, args.cpu_steps)); warm = max(0, min(args.warmup, 1))
    import torch
    workload = workload_name(args)
    if torch.cuda.is_available():
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        tb, views = make_testbed(args)
        for _ in range(args.pretrain):
            tb.train(args.batch)
        state = trained_state(tb)
        how = f"{args.pretrain} untimed steps of the CUDA Testbed (same as the own arm), read back to the host"
        del tb
        torch.cuda.empty_cache()
    else:
        from oracle import ref_api as ref
        focal = 0.5 * args.res / math.tan(0.5 * 0.6911)
        views = ref.synth_views(args.views, args.res, args.res, 4.0 * 0.33, focal)
        state = None; how = "untrained (no CUDA device to pre-train on)"; workload += "_UNTRAINED-STATE"
    r = cpu_arm(args, views, state, steps, warm, args.batch, how)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 storage / f32 accumulate",
            "data": "synthetic", "config": {"workload": workload, "global_batch": args.batch, "per_gpu_batch": args.batch,
                                            "note": "CPU restatement (oracle/) of the same training step on the host cores of rank 0; the reference has no CPU "
                                                    "implementation and its CUDA build needs the absent tiny-cuda-nn submodule",
                                            "samples_before_compaction_per_trained_sample": r["before_over_compacted"], "omp_threads": r["threads"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            "breakdown_s": r["breakdown_s"]}
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
    _lib = importlib.import_module("instant-ngp_b200._lib")
    lib = _lib.lib()
    cfg = network_config(args.log2_hashmap_size)
    tb, views = make_testbed(args)
    if world > 1:
        tb.enable_data_parallel(native=os.environ.get('NGP_DP_NATIVE', '1') == '1')
        if args.sharded_optimizer:
            tb.dp_sharded_optimizer = True      # reduce-scatter -> Adam on 1/world of the entries -> all-gather
    batch = args.batch

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clk = ClockSampler(local).__enter__()      # runs from the untimed pre-training to the end of the timed windows: every sample under the same training load
    for _ in range(args.pretrain + args.warmup):
        tb.train(batch)

    def timed(steps, upload):
        """EXACTLY `steps` steps between two CUDA events, barrier + synchronize on both sides; returns (ms max over ranks, samples
        summed over ranks, samples before compaction, launches)"""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        # the library reads each step's two sample counters back to pinned host memory (and the loss every 16th step, testbed.cu:4349) and
        # consumes them on the host before the next step starts; their running sums are read once per window instead of once per step
        n0, nb0 = tb.get_option("measured_batch_size_total"), tb.get_option("measured_batch_size_before_compaction_total")
        l0 = lib.ngp_kernel_launch_count()
        e0.record()
        for i in range(steps):
            if upload:
                tb.set_training_image(i % len(views), views[i % len(views)]["pixels"])    # H2D from pinned host memory
            tb.train(batch)
        n = int(tb.get_option("measured_batch_size_total") - n0)                           # (waits for the last step's counters)
        nb = int(tb.get_option("measured_batch_size_before_compaction_total") - nb0)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = lib.ngp_kernel_launch_count() - l0
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
            c = torch.tensor([n, nb], dtype=torch.float64, device="cuda"); dist.all_reduce(c); n, nb = int(c[0].item()), int(c[1].item())
        return ms, n, nb, launches

    def windows(repeats, upload):
        """the K-step window repeated back to back (a 20-step window is ~10 ms: one window is one clock sample and ~3 % noise);
        the reported figure is the MEDIAN window, min / max are kept beside it"""
        rows = [timed(args.steps, upload) for _ in range(repeats)]
        rows.sort(key=lambda r: r[1] / r[0])
        return rows[len(rows) // 2], rows

    ncu = os.environ.get("NGP_NCU") == "1"      # under `ncu --profile-from-start off`: capture only the timed steps
    if ncu:
        torch.cuda.profiler.start()
    t0 = time.time()
    (ms, n_samples, n_before, launches), rows = windows(1 if ncu else args.repeats, upload=False)
    t1 = time.time()
    if ncu:
        torch.cuda.profiler.stop()
    clk.__exit__()                              # (NVML polling during the upload pass was measured to slow the H2D copies: stop it first)
    clocks = clk.summary([(t0, t1)])
    time.sleep(0.25)                            # let the stopped NVML poller go away
    timed(max(args.warmup, 3), upload=True)     # untimed warm-up of the upload path (copy stream, events, first pinned-page touches)
    (ms_e2e, n_e2e, _, _), rows_e2e = windows(1 if ncu else max(3, args.repeats // 3), upload=True)
    view_bytes = int(views[0]["pixels"].numel())
    loss_steps = sum(1 for i in range(args.steps) if i % 16 == 0)
    d2h = 16 + 8.0 * loss_steps / args.steps       # two counters (+ two reserved words) every step; loss sum and density-grid mean every 16th

    # per-kernel CUDA-event timing on the launching stream over a separate pass of K steps (events add ~2 records per kernel,
    # so this pass is not the one `value` comes from)
    prefetch_hits = int(tb.prefetch_hits)
    tb.prefetch_rays = False            # kernels timed one at a time: no ray prefetch running beside them on the second stream
    lib.ngp_profile_enable(1)
    buf = C.create_string_buffer(1 << 16)
    lib.ngp_profile_report(buf, len(buf), 1)
    prof_steps = max(args.steps, 32)
    timed(prof_steps, upload=False)
    lib.ngp_profile_report(buf, len(buf), 1)
    lib.ngp_profile_enable(0)
    tb.prefetch_rays = True
    prof = json.loads(buf.value.decode())
    peaks = measured_peaks()
    traffic = profiled_traffic()
    inst = profiled_instructions()
    # roofline denominators of the hash-grid kernels measured on this device, now
    table_bytes = int(tb.n_params()) * 2
    g = C.c_double(); r1 = C.c_double(); r2 = C.c_double()
    _lib.check(lib.ngp_microbench_l2_gather(None, table_bytes, 5, C.byref(g)))
    _lib.check(lib.ngp_microbench_l2_red(None, table_bytes, 1, 5, C.byref(r1)))
    _lib.check(lib.ngp_microbench_l2_red(None, table_bytes, 2, 5, C.byref(r2)))
    enc = cfg["encoding"]; corners = enc["n_levels"] * 8
    sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
    issue_peak = 148 * 4 * sm_hz            # one warp instruction per cycle per SM sub-partition
    kernels = {}
    nested = {"train_nerf_step", "density_grid_update"}      # enclosing scopes, not kernels
    for name, r in prof.items():
        if name in nested or r["count"] == 0:
            continue
        a = algorithmic(name, cfg)
        t_launch = r["ms"] / r["count"] * 1e-3
        units = r["units"] / r["count"]
        row = {"launches_per_step": r["count"] / prof_steps, "ms_per_step": r["ms"] / prof_steps, "avg_ms": r["ms"] / r["count"], "units_per_launch": units}
        if a and a[1]:
            ach = a[1] * units / t_launch
            if a[0] == "hbm":
                row.update(bound="hbm", achieved=ach / 1e9, peak=peaks["hbm_gbs"], unit="GB/s", frac=ach / 1e9 / peaks["hbm_gbs"], per_unit=a[1])
            else:
                row.update(bound="tensor", achieved=ach / 1e12, peak=peaks["tflops_sustained"], unit="TFLOP/s", frac=ach / 1e12 / peaks["tflops_sustained"], per_unit=a[1])
            row["traffic"] = traffic.get(name)
        elif name in inst:
            # issue-bound kernels (the marcher, the compositor): warp instructions per unit from the committed ncu capture
            # (profiles/instructions.json: smsp__inst_executed.sum / units of that launch) x this launch's units / its time,
            # against one warp instruction per cycle per SM sub-partition at the SM clock sampled during the timed windows
            per_unit = inst[name]["warp_inst_per_unit"]
            ach = per_unit * units / t_launch
            row.update(bound="issue", achieved=ach / 1e9, peak=issue_peak / 1e9, unit="Gwarp-inst/s", frac=ach / issue_peak, per_unit=per_unit, traffic=traffic.get(name))
        if name in ("grid_forward", "grid_backward"):
            # the metric's own roofline ("hash-enc Gpoints/s vs L2 BW"): L2 sectors this launch requested per second — sectors per point
            # from the committed ncu capture (profiles/instructions.json; a warp's gathers coalesce when neighbouring samples share cells,
            # so sectors, not gathers, are what the memory system sees) x points / measured time — against the random-sector rate of this
            # GPU measured just now (ngp_microbench_l2_gather: one sector per gather, nothing coalesces) => <= 1 by construction
            row["gpoints_per_s"] = units / t_launch / 1e9
            spu = inst.get(name, {}).get("lts_sectors_per_unit")
            if spu:
                rate = spu * units / t_launch
                row["l2_sectors"] = {"achieved_Gsectors_per_s": rate / 1e9, "peak_Gsectors_per_s": g.value / 1e9, "frac": rate / g.value,
                                     "achieved_GB_per_s": rate * 32 / 1e9, "sectors_per_point": spu, "ncu_pct_of_peak": inst[name].get("ncu_pct_of_peak"),
                                     "peak_how": f"ngp_microbench_l2_gather over {table_bytes} B (table size): random 4-byte loads, 32 in flight per thread, 1 sector each"}
            if name == "grid_backward":
                row["l2_sectors"] = dict(row.get("l2_sectors", {}), red_f16x2_Glanes_per_s=r1.value / 1e9, red_v2_f16x2_Glanes_per_s=r2.value / 1e9,
                                         note="scattered red.global.add lanes/s of ngp_microbench_l2_red, for reference: REDs to one sector from one warp "
                                              "instruction merge, so the kernel's sector rate can exceed it")
        if name in inst and "ncu_pct_of_peak" in inst[name] and "l2_sectors" not in row:
            row["ncu_pct_of_peak"] = inst[name]["ncu_pct_of_peak"]
        kernels[name] = row
    step_ms_prof = sum(k["ms_per_step"] for k in kernels.values())
    cand = [k for k in kernels if "achieved" in kernels[k]]
    top = max(cand, key=lambda k: kernels[k]["ms_per_step"])      # the kernel group with the largest time that has a model (all main ones do)
    roof = {k: kernels[top][k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic")}
    src = {"tensor": ", sustained bf16", "hbm": ", copy bandwidth", "issue": f": 148 SMs x 4 schedulers x {sm_hz / 1e6:.0f} MHz sampled under load"}[roof["bound"]]
    roof.update(kernel=top, share_of_step=kernels[top]["ms_per_step"] / step_ms_prof, avg_ms=kernels[top]["avg_ms"],
                peak_source=(peaks["source"] if roof["bound"] != "issue" else "issue rate") + src)
    for extra in ("l2_sectors",):
        if extra in kernels[top]:
            roof[extra] = kernels[top][extra]

    rate = lambda r: r[1] / (r[0] * 1e-3)
    line = {"metric": METRIC, "value": n_samples / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 (fp32 accumulate, fp32 master weights)",
            "data": "synthetic",
            "config": {"workload": workload_name(args), "global_batch": batch * world, "per_gpu_batch": batch, "parallelism": (f"dp{world}" + ("_sharded_optimizer" if args.sharded_optimizer else "_dense_allreduce")) if world > 1 else "single",
                       "pretrain_steps_untimed": args.pretrain, "samples_before_compaction_per_step": n_before / args.steps / world,
                       "rays_per_batch": tb.stats.rays_per_batch, "n_params": int(tb.n_params()), "steps_with_prefetched_rays": prefetch_hits,
                       "l2_policy": "inputs larger than L2: 256 MB image set + 24 MB table + >100 MB sample buffers touched every step",
                       "l2_persisting_window_bytes": int(tb.get_option("l2_window_bytes"))},
            "windows": {"repeats": len(rows), "steps_each": args.steps, "value_median": rate(rows[len(rows) // 2]), "value_min": rate(rows[0]), "value_max": rate(rows[-1]),
                        "ms_per_step_min": min(r[0] for r in rows) / args.steps, "ms_per_step_max": max(r[0] for r in rows) / args.steps},
            "e2e": {"value": n_e2e / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": view_bytes, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps,
                    "api": "pyngp.Testbed.set_training_image + train (the library reads the counters back every step, the loss every 16th)", "repeats": len(rows_e2e), "value_min": rate(rows_e2e[0]), "value_max": rate(rows_e2e[-1])},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "kernels": kernels}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        state = trained_state(tb)
        r = cpu_arm(args, views, state, max(1, args.cpu_steps - 1), 1, batch, f"the own arm's parameters / occupancy grid after {int(tb.stats.training_step)} steps, read back to the host")
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_render(args):
    """--render WxH: BASELINE config 5's inference leg.  The frame is cut into `world` horizontal bands, rank r renders rows
    [r*H/world, (r+1)*H/world) with its replica of the trained parameters + bitfield (no collective on the render path); the
    bands are then gathered to rank 0 (timed separately).  ms_per_frame = max over ranks of the band time; one JSON line on rank 0."""
    import numpy as np
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    scene = importlib.import_module("instant-ngp_b200.scene")
    W, H = (int(v) for v in args.render.lower().split("x"))
    tb, views = make_testbed(args)
    if world > 1:
        tb.enable_data_parallel(native=True)
        if args.sharded_optimizer:
            tb.dp_sharded_optimizer = True
    for _ in range(args.pretrain):
        tb.train(args.batch)
    tb.sync_inference_params()
    torch.cuda.synchronize()
    # rows of rank r: every world-th row from r (default: equal work whatever the picture shows), or the contiguous band r (--render-bands)
    if args.render_bands:
        y0, y1, stride = rank * H // world, (rank + 1) * H // world, 1
    else:
        y0, y1, stride = rank, H, world
    n_rows = len(range(y0, y1, stride))
    f = scene.focal_px(H)
    band = torch.empty(n_rows, W, 4, dtype=torch.float32, device="cuda")
    cam = views[7 % len(views)]["xform"]
    def frame():
        tb.render(W, H, 1, camera_matrix=cam, focal_length=(f, f), y0=y0, y1=y1, row_stride=stride, out=band)
    for _ in range(max(args.warmup, 2)):
        frame()
    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    times = []
    for _ in range(args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync(); e0.record(); frame(); e1.record(); sync()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        times.append(ms)
    # gather of the bands to rank 0 (equal band heights when world divides H)
    gather_ms = 0.0
    if world > 1:
        max_rows = -(-H // world)
        send = band if n_rows == max_rows else torch.cat([band, band.new_zeros(max_rows - n_rows, W, 4)])
        full = [torch.empty_like(send) for _ in range(world)] if rank == 0 else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(2):
            sync(); e0.record(); dist.gather(send, full, dst=0)
            if rank == 0:      # assemble the frame: bands stack, interleaved rows alternate
                picture = (torch.cat(full) if args.render_bands else torch.stack(full, dim=1).reshape(max_rows * world, W, 4))[:H]
            e1.record(); sync()
        gather_ms = e0.elapsed_time(e1)
    times.sort()
    cov = float((band[..., 3] > 0.5).float().mean())
    if rank == 0:
        med = times[len(times) // 2]
        print(json.dumps({"metric": "nerf_render_ms_per_frame", "value": med, "unit": "ms/frame", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 2),
                          "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f16 (fp32 accumulate)", "data": "synthetic",
                          "config": {"workload": f"render_{W}x{H}_spp1_" + workload_name(args), "row_sharding": "bands" if args.render_bands else "interleaved", "ranks": world, "rows_per_rank": n_rows, "pretrain_steps_untimed": args.pretrain,
                                     "parallelism": f"tiles{world}" if world > 1 else "single"},
                          "ms_min": times[0], "ms_max": times[-1], "Mpixels_per_s": W * H / med / 1e3, "gather_to_rank0_ms": gather_ms, "gather_bytes": W * H * 16,
                          "band0_coverage": cov}), flush=True)
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
    ap.add_argument("--sharded-optimizer", dest="sharded_optimizer", action="store_true", default=True,
                    help="data parallel (default): reduce-scatter, Adam on 1/world of the entries, all-gather — measured faster than the dense all-reduce at 2 and 8 GPUs "
                         "(profiles/r02_scaling.json)")
    ap.add_argument("--dense-allreduce", dest="sharded_optimizer", action="store_false", help="data parallel: one dense all-reduce + the full optimizer on every rank")
    ap.add_argument("--render", default="", help="WxH: time the inference render, rows sharded over the ranks (BASELINE config 5), instead of training")
    ap.add_argument("--render-bands", action="store_true", help="--render: contiguous row bands per rank instead of interleaved rows")
    ap.add_argument("--repeats", type=int, default=15, help="K-step windows timed back to back; the median is reported")
    ap.add_argument("--cpu-steps", type=int, default=3, help="timed steps of the CPU arm (a full-batch step takes seconds)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.render:
        run_render(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
