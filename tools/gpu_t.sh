#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_exposure_gpu.py -m gpu -x -q -s 2>&1 | tail -40 ) > gpurun_out/t_pytest_exposure.txt
( timeout 1500 python -m pytest tests/test_nerf_gpu.py tests/test_psnr_parity_gpu.py tests/test_render_frame_gpu.py -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/t_pytest_nerf.txt
echo done
