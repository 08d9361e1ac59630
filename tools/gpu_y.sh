#!/bin/bash
# the run.py flow through the binding + the files next to what changed (loader, image mode, snapshots)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_run_script_gpu.py tests/test_nerf_loader_gpu.py tests/test_image_gpu.py tests/test_snapshot_gpu.py -x -q -m gpu > gpurun_out/y_pytest.txt 2>&1
tail -5 gpurun_out/y_pytest.txt
