"""Turn an `ncu --set full` capture of bench.py into the committed evidence (tools; needs ncu on PATH, no GPU):
profiles/<tag>_ncu_summary.txt (per-kernel metrics) and profiles/traffic.json (DRAM bytes per launch per kernel group,
read by bench.py for roofline.traffic).  usage: make_profiles.py <capture.ncu-rep> <tag> <steps captured>"""
import csv, io, json, os, subprocess, sys
rep, tag, steps = sys.argv[1], sys.argv[2], int(sys.argv[3])
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
summary = subprocess.run([sys.executable, os.path.join(root, "tools", "ncu_summary.py"), rep], capture_output=True, text=True).stdout
with open(os.path.join(root, "profiles", f"{tag}_ncu_summary.txt"), "w") as f:
    f.write(f"# ncu --set full --clock-control none, {steps} training steps of bench.py (tools/ncu_summary.py {os.path.basename(rep)})\n" + summary)
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
I = {h: i for i, h in enumerate(hdr)}
GROUP = [("grid_forward_kernel", "grid_forward"), ("grid_backward_kernel", "grid_backward"), ("adam_kernel", "adam"), ("mlp_forward_kernel", "nerf_mlp_forward"),
         ("mlp_backward_kernel", "nerf_mlp_backward"), ("fill_rollover_kernel", "fill_rollover"), ("sampler_", "sampler"), ("loss_", "loss")]
def to_bytes(v, u):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
agg = {}
for r in rows[2:]:
    name = r[I["Kernel Name"]]
    for pat, g in GROUP:
        if pat in name:
            b = sum(to_bytes(r[I[m]], units[I[m]]) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            a = agg.setdefault(g, [0.0, 0]); a[0] += b; a[1] += 1
            break
# per launch of the GROUP as bench.py counts them: one sampler / loss launch per step = all of its kernels
traffic = {g: (v[0] / steps if g in ("sampler", "loss") else v[0] / v[1]) for g, v in agg.items()}
with open(os.path.join(root, "profiles", "traffic.json"), "w") as f:
    json.dump({k: round(v) for k, v in sorted(traffic.items())}, f, indent=1)
print(json.dumps(traffic, indent=1))
