#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_psnr_parity_gpu.py -m gpu -q -s --durations=3 2>&1 | tail -30 ) > gpurun_out/e_psnr.txt
echo done
