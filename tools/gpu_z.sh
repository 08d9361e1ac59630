#!/bin/bash
mkdir -p gpurun_out
timeout 160 python -m pytest $(ls tests/test_*.py | sort | awk '$0 > "tests/test_psnr_parity_gpu.py"') -x -q -m gpu > gpurun_out/z_pytest.txt 2>&1
tail -4 gpurun_out/z_pytest.txt
