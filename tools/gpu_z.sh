#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_distortion_gpu.py -x -q -m gpu > gpurun_out/z_pytest.txt 2>&1
tail -5 gpurun_out/z_pytest.txt
