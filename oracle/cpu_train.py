"""ORACLE — TEST INFRASTRUCTURE ONLY.  One full NeRF training step on the CPU, composed of the scalar restatements
in libngp_ref.so exactly in the order Testbed::train runs them:

    training_prep_nerf -> update_density_grid_nerf      (testbed.cu:4321-4337; testbed_nerf.cu:3412-3567,4137-4152)
    generate_training_samples_nerf                       (testbed_nerf.cu:1382-1658)
    NerfNetwork::inference_mixed_precision               (nerf_network.h:116-173)
    compute_loss_kernel_train_nerf + fill_rollover       (testbed_nerf.cu:1660-2012,4061-4069)
    NerfNetwork::forward + backward (Overwrite)          (nerf_network.h:179-335)
    Trainer::optimizer_step: Ema(ExponentialDecay(Adam)) (testbed_nerf.cu:3678)
    NerfCounters::update_after_training                  (testbed_nerf.cu:3583-3609)

Users: tests/ (trajectory parity of the CUDA Testbed) and bench.py's cpu_baseline / --impl reference legs.
The product path never imports this module.
"""
import ctypes as C
import time

import numpy as np

from . import ref_api as ref

NC = 128 ** 3
LOSS_SCALE = 128.0
_LOSS = {"l2": 0, "l1": 1, "mape": 2, "smape": 3, "huber": 4, "smoothl1": 4, "logl1": 5, "relativel2": 6}


def _opt_chain(opt):
    """Ema -> ExponentialDecay -> Adam nesting (configs/nerf/base.json:5-22) flattened into one dict."""
    out = dict(ema_decay=0.0, decay_start=0, decay_interval=0, decay_base=1.0, decay_end=0xFFFFFFFF,
               learning_rate=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8, l2_reg=1e-8)
    cur = opt
    while cur is not None:
        t = cur.get("otype", "Adam").lower()
        if t == "ema":
            out["ema_decay"] = float(cur.get("decay", 0.99))
        elif t == "exponentialdecay":
            for k in ("decay_start", "decay_interval", "decay_end"):
                if k in cur:
                    out[k] = int(cur[k])
            out["decay_base"] = float(cur.get("decay_base", 1.0))
        elif t == "adam":
            for k in ("learning_rate", "beta1", "beta2", "epsilon", "l2_reg"):
                if k in cur:
                    out[k] = float(cur[k])
        cur = cur.get("nested")
    return out


class CpuNerfTrainer:
    """CPU mirror of the Nerf-mode Testbed (aabb_scale 1..128, default runtime knobs)."""

    def __init__(self, config, views, aabb_scale=1, seed=1337):
        self.cfg = ref.nerf_cfg_from_json(config)
        self.images, self._keep = ref.images_from_views(views)
        self.seed = seed
        self.opt = _opt_chain(config.get("optimizer", {"otype": "Adam"}))
        self.loss_type = _LOSS[config.get("loss", {}).get("otype", "L2").lower()]
        n, _, _, off_grid = ref.nerf_n_params(self.cfg)
        self.n_params, self.n_matrix = n, off_grid
        self.master = ref.nerf_init(self.cfg, seed)
        self.params = self.master.astype(np.float16)
        self.m = np.zeros(n, np.float32); self.v = np.zeros(n, np.float32); self.steps = np.zeros(n, np.uint32)
        self.ema = np.zeros(n, np.float32); self.ema_h = self.params.copy()
        self.lr = self.opt["learning_rate"]; self.opt_steps = 0
        # load_nerf_post (testbed_nerf.cu:3046-3112)
        half = 0.5 * min(128, aabb_scale)
        self.aabb = ref.make_aabb((0.5 - half,) * 3, (0.5 + half,) * 3)
        self.max_cascade = int(np.ceil(np.log2(aabb_scale))) if aabb_scale > 1 else 0
        self.cone = 0.0 if aabb_scale <= 1 else 1.0 / 256.0
        ncell = NC * (self.max_cascade + 1)
        self.grid = np.zeros(ncell, np.float32)
        self.bitfield = np.zeros(NC * 8 // 8, np.uint8)
        self.mean_density = 0.0
        # reset_network (testbed.cu:3906-3919)
        s = np.zeros(1, np.uint32)
        ref.lib().ref_pcg32_fill_uint(ref.u64(seed), ref.u64(0), ref.P(s), C.c_size_t(1))
        self.dg_seed, self.dg_skip, self.dg_step = int(s[0]), 0, 0
        self.rng_skip = 1
        self.training_step = 0
        self.rays_per_batch = 1 << 12
        self.measured = 0; self.measured_before = 0
        self.loss = 0.0
        self.timings = {}

    def adopt_state(self, st):
        """Continue from a trained state (dict of numpy arrays / ints, e.g. read back from the CUDA Testbed): parameters, optimizer
        moments, occupancy grid + bitfield, step counters.  bench.py uses it so that the CPU arm times the SAME steady-state step
        as the GPU arm (sparsified occupancy grid, settled rays_per_batch) without training for minutes on the host first."""
        self.master = np.ascontiguousarray(st["master"], np.float32); self.params = np.ascontiguousarray(st["params"], np.float16)
        assert self.master.size == self.n_params and self.params.size == self.n_params
        for k in ("m", "v", "ema"):
            if k in st: setattr(self, k, np.ascontiguousarray(st[k], np.float32))
        if "steps" in st: self.steps = np.ascontiguousarray(st["steps"], np.uint32)
        if "ema_h" in st: self.ema_h = np.ascontiguousarray(st["ema_h"], np.float16)
        self.grid = np.ascontiguousarray(st["density_grid"], np.float32); self.bitfield = np.ascontiguousarray(st["bitfield"], np.uint8)
        self.mean_density = float(st["mean_density"])
        self.training_step = int(st["training_step"]); self.opt_steps = int(st.get("opt_steps", self.training_step))
        self.rays_per_batch = int(st["rays_per_batch"]); self.measured_before = int(st["measured_before"]); self.measured = int(st.get("measured", 0))
        self.rng_skip = 1 + self.training_step * (1 << 32)          # m_rng.advance() once per step (testbed_nerf.cu:4127)
        # occupancy-grid refreshes so far: every clamp(step/16, 1, 16) steps (testbed.cu:4321); two rng advances per refresh
        n_upd = sum(1 for t in range(self.training_step) if t % min(max(t // 16, 1), 16) == 0)
        self.dg_step = int(st.get("dg_step", n_upd)); self.dg_skip = int(st.get("dg_skip", n_upd * 2 * (1 << 32)))

    def _t(self, name, t0):
        self.timings[name] = self.timings.get(name, 0.0) + time.perf_counter() - t0

    # update_density_grid_nerf (testbed_nerf.cu:3412-3536)
    def update_density_grid(self, n_uniform, n_nonuniform, decay=0.95):
        t0 = time.perf_counter()
        l = ref.lib()
        ncas = self.max_cascade + 1
        if self.training_step == 0:
            self.dg_step = 0
            ref.mark_untrained(self.grid, self.images)
        tmp = np.zeros_like(self.grid)
        pos_u, idx_u = ref.generate_grid_samples(n_uniform, self.dg_seed, self.dg_skip, self.dg_step, self.aabb, self.grid, ncas, -0.01)
        self.dg_skip += 1 << 32
        if n_nonuniform:
            pos_n, idx_n = ref.generate_grid_samples(n_nonuniform, self.dg_seed, self.dg_skip, self.dg_step, self.aabb, self.grid, ncas, 0.01)
            pos = np.concatenate([pos_u, pos_n]); idx = np.concatenate([idx_u, idx_n])
        else:
            pos, idx = pos_u, idx_u
        self.dg_skip += 1 << 32
        dens = ref.nerf_density(self.cfg, self.params, pos)
        l.ref_splat_grid_samples(ref.u32(idx.size), ref.P(idx), ref.P(dens), ref.P(tmp), ref.i32(3))
        l.ref_ema_grid_samples(ref.u32(self.grid.size), ref.f32(decay), ref.P(self.grid), ref.P(tmp))
        self.dg_step += 1
        self.bitfield, self.mean_density = ref.update_bitfield(self.grid, ncas)
        self._t("density_grid", t0)

    def train(self, batch):
        """One Testbed::train(batch) step.  Returns the number of (compacted) samples trained on."""
        l = ref.lib()
        skip_n = min(max(self.training_step // 16, 1), 16)
        if self.training_step % skip_n == 0:
            ncas = self.max_cascade + 1
            if self.training_step < 256:
                self.update_density_grid(NC * ncas, 0)
            else:
                self.update_density_grid(NC // 4 * ncas, NC // 4 * ncas)
        get_loss = self.training_step % 16 == 0
        R = self.rays_per_batch
        max_samples = batch * 16
        if self.measured_before == 0:
            self.measured_before = max_inf = max_samples
        else:
            max_inf = (min(self.measured_before, max_samples) + 255) // 256 * 256
        t0 = time.perf_counter()
        o = ref.generate_training_samples(R, self.aabb, max_inf, self.seed, self.rng_skip, self.images, self.bitfield, self.max_cascade, True, self.cone)
        self._t("sampler", t0); t0 = time.perf_counter()
        n_written = int(min(o["numsteps_counter"], max_inf))
        n_inf = (n_written + 255) // 256 * 256
        coords = o["coords"]
        out = np.zeros((max_inf, 4), np.float16)
        out[:n_inf] = ref.nerf_forward(self.cfg, self.params, coords[:n_inf])
        self._t("inference", t0); t0 = time.perf_counter()
        lc = ref.LossCfg(self.loss_type, 2, 3, 1, 0, 1, (ref.f32 * 3)(0, 0, 0), LOSS_SCALE, 0.1, self.mean_density)
        lo = ref.compute_loss(R, self.aabb, self.seed, self.rng_skip, batch, o["ray_counter"], lc, self.images, out, o["ray_indices"], o["rays"],
                              o["numsteps"], coords)
        nc = int(min(lo["compacted_counter"], batch))
        if o["numsteps_counter"] == 0 or nc == 0:
            raise RuntimeError("Nerf training generated 0 samples. Aborting training.")
        dl, cc = lo["dloss"], lo["coords_compacted"]
        l.ref_fill_rollover_half(ref.u32(batch), ref.u32(4), ref.u32(nc), ref.P(dl), ref.i32(1))
        l.ref_fill_rollover_float(ref.u32(batch), ref.u32(7), ref.u32(nc), ref.P(cc))
        self._t("loss", t0); t0 = time.perf_counter()
        g = ref.nerf_backward(self.cfg, self.params, cc, dl)
        grads = g.astype(np.float32).astype(np.float16)
        self._t("forward_backward", t0); t0 = time.perf_counter()
        ac = ref.AdamCfg(self.lr, self.opt["beta1"], self.opt["beta2"], self.opt["epsilon"], self.opt["l2_reg"], LOSS_SCALE, self.n_matrix, 1, 1)
        ref.adam_step(ac, self.master, self.params, grads, self.m, self.v, self.steps)
        self.opt_steps += 1
        if self.opt["ema_decay"] > 0:
            l.ref_ema_step(ref.u64(self.n_params), ref.f32(self.opt["ema_decay"]), ref.u32(self.opt_steps), ref.P(self.params), ref.P(self.ema), ref.P(self.ema_h))
        else:
            self.ema_h = self.params
        sc = self.opt_steps
        if self.opt["decay_interval"] > 0 and self.opt["decay_start"] <= sc < self.opt["decay_end"] and (sc - self.opt["decay_start"]) % self.opt["decay_interval"] == 0:
            self.lr *= self.opt["decay_base"]
        self._t("optimizer", t0)
        # NerfCounters::update_after_training
        self.training_step += 1
        self.rng_skip += 1 << 32
        self.measured_before = int(o["numsteps_counter"]); self.measured = int(lo["compacted_counter"])
        if get_loss:
            self.loss = float(lo["loss"][:R].sum(dtype=np.float64)) * self.measured / batch
        self.rays_per_batch = int(l.ref_next_rays_per_batch(ref.u32(R), ref.u32(batch), ref.u32(self.measured)))
        return nc
