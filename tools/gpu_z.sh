#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_snapshot_gpu.py tests/test_depth_gpu.py tests/test_max_level_gpu.py tests/test_exposure_gpu.py tests/test_extrinsics_gpu.py -x -q -m gpu > gpurun_out/z_pytest.txt 2>&1
tail -5 gpurun_out/z_pytest.txt
timeout 1500 python bench.py --no-cpu-baseline > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err
tail -c 300 gpurun_out/z_bench.json
