#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_nerf_gpu.py tests/test_run_script_gpu.py -x -q -m gpu -k "deferred or control_law or run_py or set_training_image" > gpurun_out/z_pytest.txt 2>&1
tail -5 gpurun_out/z_pytest.txt
