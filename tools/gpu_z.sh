#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_extrinsics_gpu.py -q -m gpu -s > gpurun_out/z_pytest.txt 2>&1
tail -5 gpurun_out/z_pytest.txt
