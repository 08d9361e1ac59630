"""Host time per training step of the upload path (tools, not product): how long does the host need to ENQUEUE a step
(set_training_image + train + stats) when it never waits for the GPU?  Compared with the GPU's own time per step it says whether
bench.py's e2e figure is host- or device-bound."""
import argparse, importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

args = argparse.Namespace(views=100, res=800, log2_hashmap_size=19)
tb, views = bench.make_testbed(args)
B = 1 << 18
for _ in range(600):
    tb.train(B)
torch.cuda.synchronize()
pyngp = importlib.import_module("instant-ngp_b200.pyngp")


def loop(n, upload, parts):
    t_up = t_tr = t_st = 0.0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        a = time.perf_counter()
        if upload:
            tb.set_training_image(i % len(views), views[i % len(views)]["pixels"])
        b = time.perf_counter()
        tb.train(B)
        c = time.perf_counter()
        if parts:
            st = tb.stats; _ = st.loss
        d = time.perf_counter()
        t_up += b - a; t_tr += c - b; t_st += d - c
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"upload={upload} stats_every_step={parts} n={n}: host loop {1e3 * (t1 - t0) / n:.3f} ms/step (set_training_image {1e3 * t_up / n:.3f}, train {1e3 * t_tr / n:.3f}, stats {1e3 * t_st / n:.3f}); "
          f"with the final sync {1e3 * (t2 - t0) / n:.3f} ms/step")


for n in (20, 40):
    loop(n, False, True)
    loop(n, True, True)
    loop(n, False, False)
    loop(n, True, False)
