#!/bin/bash
# scratch wrapper for one gpurun call: edit the test selection, run `bash tools/gpurun_retry.sh gpurun_out/z_call.log 1800 'bash tools/gpu_z.sh'`
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_nerf_gpu.py -x -q -m gpu > gpurun_out/z_pytest.txt 2>&1
tail -5 gpurun_out/z_pytest.txt
