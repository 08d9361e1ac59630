#!/bin/bash
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_psnr_parity_gpu.py -x -q -m gpu -s 2>&1 | grep "PSNR:\|passed\|failed\|Error" | cut -c1-140 ) > gpurun_out/z_pytest.txt 2>&1
cat gpurun_out/z_pytest.txt
