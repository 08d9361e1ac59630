"""Turn an `ncu --set full` capture of bench.py into the committed evidence (tools; needs ncu on PATH, no GPU):
profiles/<tag>_ncu_summary.txt (per-kernel metrics) and profiles/traffic.json (DRAM bytes per launch per kernel group,
read by bench.py for roofline.traffic) and profiles/instructions.json (warp instructions per ray of the issue-bound kernel groups, read by
bench.py for their issue roofline).  usage: make_profiles.py <capture.ncu-rep> <tag> <steps captured> [rays per step of the captured run]"""
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
GROUP = [("grid_forward_kernel", "grid_forward"), ("grid_backward_kernel", "grid_backward"), ("adam_kernel", "adam"), ("mlp_forward", "nerf_mlp_forward"),
         ("mlp_backward", "nerf_mlp_backward"), ("fill_rollover_kernel", "fill_rollover"), ("sampler_", "sampler"), ("loss_", "loss")]
def to_bytes(v, u):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
agg = {}
for r in rows[2:]:
    name = r[I["Kernel Name"]]
    for pat, g in GROUP:
        if pat in name:
            b = sum(to_bytes(r[I[m]], units[I[m]]) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            a = agg.setdefault(g, [0.0, 0, 0.0, 0.0, {}]); a[0] += b; a[1] += 1
            def num(m):
                return float(r[I[m]].replace(",", "")) if m in I and r[I[m]] not in ("", "n/a") else 0.0
            a[3] += num("lts__t_sectors.sum")
            for m in ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
                      "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active", "gpu__time_duration.sum"):
                a[4].setdefault(m, []).append(num(m))
            a[2] += float(r[I["smsp__inst_executed.sum"]].replace(",", "")) if "smsp__inst_executed.sum" in I and r[I["smsp__inst_executed.sum"]] not in ("", "n/a") else 0.0
            break
# per launch of the GROUP as bench.py counts them: one sampler / loss launch per step = all of its kernels
traffic = {g: (v[0] / steps if g in ("sampler", "loss") else v[0] / v[1]) for g, v in agg.items()}
with open(os.path.join(root, "profiles", "traffic.json"), "w") as f:
    json.dump({k: round(v) for k, v in sorted(traffic.items())}, f, indent=1)
print(json.dumps(traffic, indent=1))

if len(sys.argv) > 4:
    rays = float(sys.argv[4])
    inst = {g: {"warp_inst_per_unit": agg[g][2] / steps / rays, "unit": "ray", "warp_inst_per_step": agg[g][2] / steps, "rays_per_step_of_capture": rays,
                "source": f"profiles/{tag}_ncu_summary.txt (smsp__inst_executed.sum over the group's kernels of one step)"} for g in ("sampler", "loss") if g in agg}
    # hash-grid kernels: L2 sectors per point of the captured launches (units = points; grid.x = ceil(points / 256) blocks) and ncu's own
    # pct-of-peak figures; bench.py turns sectors per point into a live sector rate against the measured random-sector peak
    for g, pat in (("grid_forward", "grid_forward_kernel"), ("grid_backward", "grid_backward_kernel")):
        pts = 0.0
        for r in rows[2:]:
            if pat in r[I["Kernel Name"]]:
                pts += float(r[I["launch__grid_size"]].replace(",", "")) / 4.0 * 256.0        # grid = (ceil(n / 256), 4 chunks)
        if g in agg and pts:
            inst[g] = {"lts_sectors_per_unit": agg[g][3] / pts, "unit": "point (upper bound on points: whole 256-point blocks)",
                       "ncu_pct_of_peak": {k.split(".")[0]: round(sum(v) / len(v), 2) for k, v in agg[g][4].items() if "duration" not in k},
                       "ncu_avg_us": round(sum(agg[g][4]["gpu__time_duration.sum"]) / len(agg[g][4]["gpu__time_duration.sum"]), 2),
                       "source": f"profiles/{tag}_ncu_summary.txt"}
    for g in ("nerf_mlp_forward", "nerf_mlp_backward", "adam"):
        if g in agg:
            inst[g] = {"ncu_pct_of_peak": {k.split(".")[0]: round(sum(v) / len(v), 2) for k, v in agg[g][4].items() if "duration" not in k},
                       "ncu_avg_us": round(sum(agg[g][4]["gpu__time_duration.sum"]) / len(agg[g][4]["gpu__time_duration.sum"]), 2), "source": f"profiles/{tag}_ncu_summary.txt"}
    with open(os.path.join(root, "profiles", "instructions.json"), "w") as f:
        json.dump(inst, f, indent=1)
    print(json.dumps(inst, indent=1))
