#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_nerf_gpu.py tests/test_exposure_gpu.py -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/v_pytest.txt
timeout 900 python tools/bench_fox.py --steps 35000 --profile 64 --eval-views 10 > gpurun_out/v_fox_35k.json 2> gpurun_out/v_fox_35k.err
echo done
