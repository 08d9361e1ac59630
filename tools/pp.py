"""compact view of tools/phase_profile.py lines read from stdin"""
import sys, json
for l in sys.stdin:
    if not l.startswith('{"step"'): continue
    d = json.loads(l); k = d["kernels_ms_per_step"]
    print(d["ms_per_step_wall"], d["measured"], d["before"], " ".join(f"{n.replace('nerf_','')}={v:.3f}" for n, v in k.items() if n != "train_nerf_step"))
