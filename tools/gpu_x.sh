#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 2400 python -m pytest tests/test_render_frame_gpu.py tests/test_sdf_gpu.py tests/test_snapshot_gpu.py tests/test_zz_no_alloc.py tests/test_psnr_parity_gpu.py -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/x_pytest_rest.txt
echo done
