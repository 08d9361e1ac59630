#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_nerf_gpu.py -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/r_pytest_nerf.txt
timeout 900 python tools/bench_fox.py --steps 5000 --profile 64 --eval-views 4 > gpurun_out/r_fox_5k_walkahead.json 2> gpurun_out/r_fox_5k_walkahead.err
echo done
