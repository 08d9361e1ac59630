#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q -k "baseline_sizes or headline_shape or opencv_lens or psnr or albert" --durations=10 -s 2>&1 | tail -60 ) > gpurun_out/d_pytest_new.txt
echo done
