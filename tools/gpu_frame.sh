#!/bin/bash
# round 2: boundary tests (render with the testbed's view state, output stage, run.py evaluation loop, Halton / Sobol image samplers)
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_render_frame_gpu.py tests/test_image_gpu.py -m gpu -x -q 2>&1 | tail -40 ) > gpurun_out/k_pytest_frame.txt
( timeout 1200 python -m pytest tests/test_nerf_gpu.py tests/test_snapshot_gpu.py tests/test_nerf_loader_gpu.py -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/k_pytest_nerf.txt
echo done
