#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_exposure_gpu.py -m gpu -x -q -s 2>&1 | tail -25 ) > gpurun_out/w_pytest.txt
echo done
