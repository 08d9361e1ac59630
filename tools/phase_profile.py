"""Per-kernel step profile at several points of a NeRF training run (tools, not product): how the step mix moves as the
occupancy grid sparsifies and rays_per_batch adapts."""
import ctypes as C, importlib, json, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
pyngp = importlib.import_module("instant-ngp_b200.pyngp"); scene = importlib.import_module("instant-ngp_b200.scene")
lib = importlib.import_module("instant-ngp_b200._lib").lib()
nv, res = int(sys.argv[1]) if len(sys.argv) > 1 else 100, int(sys.argv[2]) if len(sys.argv) > 2 else 800
views = scene.make_views(nv, res, res)
tb = pyngp.Testbed(pyngp.TestbedMode.Nerf); tb.set_nerf_dataset(views, aabb_scale=1); tb.reload_network_from_json(bench.network_config(19))
tb.sampler_lanes_per_ray = int(os.environ.get('NGP_LANES', '0')); tb.prefetch_rays = os.environ.get('NGP_PREFETCH', '1') == '1'
buf = C.create_string_buffer(1 << 16)
def psnr(view):
    v = views[view]; W = 200; f = scene.focal_px(W)
    img = tb.render(W, W, spp=1, camera_matrix=v["xform"], focal_length=(f, f))
    px = v["pixels"].numpy()[::res // W, ::res // W].astype(np.float32) / 255.0
    a = px[..., 3:4]; lin = np.where(px[..., :3] <= 0.04045, px[..., :3] / 12.92, ((px[..., :3] + 0.055) / 1.055) ** 2.4) * a
    return -10 * np.log10(max(float(((img[..., :3] - lin) ** 2).mean()), 1e-10))
marks = [int(x) for x in (sys.argv[3].split(",") if len(sys.argv) > 3 else "64,256,512,1000,2000,4000".split(","))]
step = 0; NSTEP = int(os.environ.get("NGP_NSTEP", "32"))
if os.environ.get('NGP_MAIN_PRIORITY'):      # run the training stream above the prefetch stream
    _hp = torch.cuda.Stream(priority=int(os.environ['NGP_MAIN_PRIORITY'])); torch.cuda.set_stream(_hp)
for m in marks:
    while step < m: tb.train(1 << 18); step += 1
    lib.ngp_profile_enable(1); lib.ngp_profile_report(buf, len(buf), 1)
    if os.environ.get('NGP_NCU') == '1': torch.cuda.profiler.start()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(NSTEP): tb.train(1 << 18); step += 1
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / NSTEP * 1e3
    lib.ngp_profile_report(buf, len(buf), 1); lib.ngp_profile_enable(0)
    prof = json.loads(buf.value.decode()); st = tb.stats
    print(json.dumps({"step": step, "ms_per_step_wall": round(dt, 3), "rays_per_batch": st.rays_per_batch, "measured": st.measured_batch_size,
                      "before": st.measured_batch_size_before_compaction, "loss": st.loss, "psnr_view0": round(psnr(0), 2),
                      "kernels_ms_per_step": {k: round(v["ms"] / NSTEP, 4) for k, v in prof.items()}}), flush=True)
if os.environ.get('NGP_DUMP'):      # state of the last sampler launch, for offline analysis of the walk (tools/walk_stats.py)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32); nr = int(c[0])
    np.savez_compressed(os.environ['NGP_DUMP'], counters=c, bitfield=tb.density_grid(1).cpu().numpy(),
                        rays=tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6),
                        numsteps=tb.debug_tensor("numsteps", torch.int32, nr * 2).cpu().numpy().reshape(nr, 2))
