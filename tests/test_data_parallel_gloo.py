"""Data-parallel host logic on CPU (gloo, world_size 2): rank r marches the global rays [r*R, (r+1)*R) of a 2R-ray step
(ref_set_shard == ngp_testbed_set_shard), the loss is normalised by the global ray count, and one all_reduce(sum) over
the flat gradient vector must reproduce the single-process gradient of the 2R-ray step.  Also checks the lock-step
rays_per_batch rule that pyngp.Testbed.train() applies to the all-reduced sample counters."""
import math
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import json
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = {
    "loss": {"otype": "Huber"},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 2, "log2_hashmap_size": 12, "base_resolution": 8, "per_level_scale": 1.5},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
}
R, MAX_SAMPLES, BATCH = 96, 1 << 17, 1 << 12


def _gradient(ref, views, ray_offset, n_rays, n_global):
    """sampler -> network -> loss -> backward on rays [ray_offset, ray_offset + n_rays) of an n_global-ray step; fp64 gradient"""
    images, keep = ref.images_from_views(views)
    cfg = ref.nerf_cfg_from_json(CFG)
    params = ref.nerf_init(cfg, 1337).astype(np.float16)
    params[: 64 * 16 + 16 * 64] *= 20          # make the network output non-trivial
    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    bitfield = np.full(128 ** 3 * 8 // 8, 0xFF, np.uint8)
    ref.lib().ref_set_shard(ref.u32(ray_offset), ref.u32(n_global))
    try:
        o = ref.generate_training_samples(n_rays, aabb, MAX_SAMPLES, 1337, 1, images, bitfield, 0, True, 0.0)
        n = int(min(o["numsteps_counter"], MAX_SAMPLES)); n_pad = (n + 255) // 256 * 256
        out = np.zeros((MAX_SAMPLES, 4), np.float16)
        out[:n_pad] = ref.nerf_forward(cfg, params, o["coords"][:n_pad])
        lc = ref.LossCfg(4, 2, 3, 1, 0, 1, (ref.f32 * 3)(0, 0, 0), 128.0, 0.1, 1.0)
        lo = ref.compute_loss(n_rays, aabb, 1337, 1, MAX_SAMPLES, o["ray_counter"], lc, images, out, o["ray_indices"], o["rays"], o["numsteps"], o["coords"])
        nc = int(min(lo["compacted_counter"], MAX_SAMPLES)); nc_pad = (nc + 255) // 256 * 256
        dl = lo["dloss"][:nc_pad].copy(); dl[nc:] = 0
        g = ref.nerf_backward(cfg, params, lo["coords_compacted"][:nc_pad].copy(), dl)
    finally:
        ref.lib().ref_set_shard(ref.u32(0), ref.u32(0))
    return g, int(o["numsteps_counter"]), int(lo["compacted_counter"]), float(lo["loss"][:n_rays].sum(dtype=np.float64))


def _worker(rank, world, port):
    sys.path.insert(0, ROOT)
    from oracle import ref_api as ref
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    f = 0.5 * 48 / math.tan(0.5 * 0.6911)
    views = ref.synth_views(4, 48, 48, 4.0 * 0.33, f)
    g, before, after, loss = _gradient(ref, views, rank * R, R, world * R)
    t = torch.from_numpy(g); dist.all_reduce(t)                                   # the per-step gradient all-reduce
    c = torch.tensor([after, before], dtype=torch.int64); dist.all_reduce(c)      # lock-step counters (pyngp.Testbed.train)
    l = torch.tensor([loss], dtype=torch.float64); dist.all_reduce(l)
    if rank == 0:
        g_all, before_all, after_all, loss_all = _gradient(ref, views, 0, world * R, world * R)
        print("RESULT " + json.dumps(dict(err=float(np.abs(t.numpy() - g_all).max()), scale=float(np.abs(g_all).max()), counters=[int(c[0]), int(c[1])],
                                          single=[after_all, before_all], loss=[float(l[0]), loss_all],
                                          rpb=int(ref.lib().ref_next_rays_per_batch(ref.u32(R), ref.u32(BATCH), ref.u32(max(int(c[0]) // world, 1)))))), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_gradients_sum_to_the_global_step():
    port = 29500 + os.getpid() % 1000
    env = dict(os.environ, OMP_NUM_THREADS="4")
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", str(r), "2", str(port)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                              text=True, env=env) for r in range(2)]
    outs = []
    for p in procs:
        out, _ = p.communicate(timeout=300)
        outs.append(out)
        assert p.returncode == 0, out[-2000:]
    line = [l for l in outs[0].splitlines() if l.startswith("RESULT ")]
    assert line, outs[0][-2000:]
    res = json.loads(line[0][7:])
    assert res["scale"] > 0
    assert res["err"] <= 1e-9 * res["scale"] + 1e-12, res          # fp64 sums of the same per-sample terms, different order
    assert res["counters"] == res["single"], res                   # the two shards hold exactly the global step's samples
    assert abs(res["loss"][0] - res["loss"][1]) <= 1e-6 * abs(res["loss"][1]) + 1e-12
    assert res["rpb"] % 256 == 0 and res["rpb"] >= 256


def _worker_sharded(rank, world, port):
    """sharded optimizer (csrc/testbed.cu sharded_step / Trainer::optimizer_step_shard): reduce-scatter of the padded fp16 gradient vector,
    Adam + EMA on this rank's slice with the matrix-parameter boundary moved into slice coordinates, all-gather of the fp16 parameters
    == all-reduce + the full optimizer on every rank, entry for entry"""
    sys.path.insert(0, ROOT)
    from oracle import ref_api as ref
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n, nm = 50000 + 8 * 3, 7168
    n_cap = (n + 2047) // 2048 * 2048; shard = n_cap // world
    assert shard % 128 == 0
    rng = np.random.default_rng(5)                                   # same initial state on every rank
    master = rng.normal(0, 0.1, n).astype(np.float32); params = master.astype(np.float16)
    m = np.zeros(n, np.float32); v = np.zeros(n, np.float32); steps = np.zeros(n, np.uint32)
    full = [a.copy() for a in (master, params, m, v, steps)]
    params_cap = np.zeros(n_cap, np.float16); params_cap[:n] = params
    for step in range(3):
        local = np.random.default_rng(100 + 10 * step + rank)       # this rank's gradient
        g = np.zeros(n_cap, np.float32)
        g[:n] = local.normal(0, 1, n) * (local.uniform(size=n) < 0.3)
        # --- dense path: all-reduce, full Adam (what every rank does without the option) ---
        t = torch.from_numpy(g.copy()); dist.all_reduce(t)
        g_sum = t.numpy().astype(np.float16)
        ref.adam_step(ref.AdamCfg(1e-2, 0.9, 0.99, 1e-15, 1e-6, 128.0, nm, 1, 1), full[0], full[1], g_sum[:n].copy(), full[2], full[3], full[4])
        # --- sharded path: reduce each slice to its owner, Adam on [start, end), all-gather ---
        mine = None
        for r in range(world):
            sl = torch.from_numpy(g[r * shard:(r + 1) * shard].copy()); dist.reduce(sl, dst=r)
            if r == rank:
                mine = sl.numpy().astype(np.float16)
        start = rank * shard; end = min(start + shard, n)
        nm_local = max(0, min(nm - start, end - start)) if nm > start else 0
        if end > start:
            sm, sp, smm, sv, ss = master[start:end].copy(), params_cap[start:end].copy(), m[start:end].copy(), v[start:end].copy(), steps[start:end].copy()
            ref.adam_step(ref.AdamCfg(1e-2, 0.9, 0.99, 1e-15, 1e-6, 128.0, nm_local, 1, 1), sm, sp, mine[:end - start].copy(), smm, sv, ss)
            master[start:end], params_cap[start:end], m[start:end], v[start:end], steps[start:end] = sm, sp, smm, sv, ss
        parts = [torch.zeros(shard, dtype=torch.float16) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(params_cap[start:start + shard].copy()))
        params_cap = torch.cat(parts).numpy()
    ok = bool(np.array_equal(params_cap[:n].view(np.uint16), full[1].view(np.uint16))) and bool(np.array_equal(steps[start:end], full[4][start:end])) \
        and bool(np.array_equal(master[start:end].view(np.uint32), full[0][start:end].view(np.uint32)))
    flag = torch.tensor([int(ok)]); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("RESULT " + json.dumps(dict(ok=int(flag[0]), shard=shard, n_cap=n_cap)), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_optimizer_equals_allreduce_plus_full_adam():
    port = 30600 + os.getpid() % 1000
    env = dict(os.environ, OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker-sharded", str(r), "2", str(port)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True, env=env) for r in range(2)]
    outs = []
    for p in procs:
        out, _ = p.communicate(timeout=300)
        outs.append(out)
        assert p.returncode == 0, out[-2000:]
    line = [l for l in outs[0].splitlines() if l.startswith("RESULT ")]
    assert line, outs[0][-2000:]
    res = json.loads(line[0][7:])
    assert res["ok"] == 1 and res["shard"] * 2 == res["n_cap"], res


if __name__ == "__main__" and len(sys.argv) >= 5 and sys.argv[1] == "--worker":
    _worker(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]))
if __name__ == "__main__" and len(sys.argv) >= 5 and sys.argv[1] == "--worker-sharded":
    _worker_sharded(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]))
