#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_nerf_gpu.py -m gpu -x -q 2>&1 | tail -12 ) > gpurun_out/j_pytest_nerf.txt
timeout 300 python tools/timeline.py 4 > gpurun_out/j_timeline_loss.txt 2>&1
NGP_PREFETCH_AFTER=mlp timeout 300 python tools/timeline.py 4 > gpurun_out/j_timeline_mlp.txt 2>&1
for v in loss mlp; do NGP_PREFETCH_AFTER=$v timeout 600 python bench.py --steps 64 --warmup 16 --repeats 9 --no-cpu-baseline > gpurun_out/j_bench_$v.json 2> gpurun_out/j_bench_$v.err; done
echo done
