"""Counts of Blackwell-specific SASS mnemonics in the headline kernels of libngp_b200.so (tools; cuobjdump + c++filt)."""
import collections, re, subprocess, sys, os
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sass = subprocess.run(["cuobjdump", "-sass", os.path.join(root, "instant-ngp_b200", "libngp_b200.so")], capture_output=True, text=True).stdout
WANT = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTCBAR", "SYNCS", "REDG.E.ADD.F16x2", "REDG.E.ADD.F16x4", "REDG.E.ADD.F32x4", "REDG.E.ADD.F32", "RED.E.ADD.F16x2", "ATOMG", "LDG.E.64", "LDG.E.128", "STG.E.128"]
HEAD = ["mlp_forward_ts_kernelILj32ELj1ELj2ELi0E", "mlp_forward_kernelILj32ELj1ELj2ELi0E", "mlp_backward_kernelILj32ELj1ELj2ELi0ELb1ELb0ELb0E", "mlp_backward_kernelILj32ELj1ELj2ELi0ELb1ELb1ELb0E", "grid_forward_kernelILj3ELj2E", "grid_backward_kernelILj3ELj2E", "adam_kernel",
        "sampler_count_group_kernelILj8ELb1ELj1E", "sampler_expand_kernel", "loss_forward_group_kernelILj32E", "loss_backward_group_kernelILj32E"]
cur, counts, sample = None, collections.defaultdict(collections.Counter), {}
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = next((h for h in HEAD if h in m.group(1)), None); continue
    if cur is None: continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
    if not m: continue
    op = m.group(1)
    for w in WANT:
        if op.startswith(w):
            counts[cur][w] += 1; sample.setdefault(w, line.strip()[:110]); break
print("# cuobjdump -sass instant-ngp_b200/libngp_b200.so (sm_100a): Blackwell-specific mnemonics in the headline kernels")
for h in HEAD:
    name = subprocess.run(["c++filt", "_ZN3ngp" + h] , capture_output=True, text=True).stdout.strip()
    print(f"{h:45s} " + "  ".join(f"{k}={v}" for k, v in sorted(counts[h].items())))
print("\n# one sample line per mnemonic")
for w, l in sample.items(): print(f"{w:18s} {l}")
