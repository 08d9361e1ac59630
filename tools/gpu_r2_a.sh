#!/bin/bash
# round-2 GPU call A: parity suite, tensor-core / L2 micro-benchmarks, A/B of the L2 window and the 16-byte RED merge
mkdir -p gpurun_out && cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/a_gpu.txt 2>&1
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/a_pytest.txt
( timeout 120 tools/ubench_tc 148; timeout 60 tools/ubench_tc 1 ) > gpurun_out/a_ubench_tc.txt 2>&1
timeout 300 python - > gpurun_out/a_l2.txt 2>&1 <<'PY'
import ctypes as C, importlib, json, torch
torch.cuda.set_device(0); torch.zeros(1, device="cuda")
lib = importlib.import_module("instant-ngp_b200._lib").lib()
out = {}
for name, mb in (("24.4MB", 24.4e6), ("96MB", 96e6), ("800MB", 800e6)):
    v = C.c_double(); assert lib.ngp_microbench_l2_gather(None, int(mb), 9, C.byref(v)) == 0; out[f"gather_{name}_G_per_s"] = v.value / 1e9
    for vec in (1, 2, 4):
        assert lib.ngp_microbench_l2_red(None, int(mb), vec, 9, C.byref(v)) == 0; out[f"red_v{vec}_{name}_Glanes_per_s"] = v.value / 1e9
print(json.dumps(out, indent=1))
PY
for cfg in "0 0" "1 0" "0 1" "1 1"; do set -- $cfg
  NGP_DEBUG=1 NGP_L2_PERSIST=$1 NGP_RED_V4=$2 timeout 600 python bench.py --steps 64 --warmup 16 --no-cpu-baseline > gpurun_out/a_bench_l2$1_v4$2.json 2> gpurun_out/a_bench_l2$1_v4$2.err
done
echo done
