#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; : > gpurun_out/k_flaky.txt
for i in 1 2 3 4; do ( timeout 200 python -m pytest tests/test_nerf_gpu.py -m gpu -x -q -k "camera_setters" 2>&1 | tail -2 ) >> gpurun_out/k_flaky.txt; done
( timeout 900 python -m pytest tests/test_nerf_gpu.py -m gpu -q 2>&1 | tail -6 ) >> gpurun_out/k_flaky.txt
git stash -q 2>/dev/null
