#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
for lanes in 2 4 8 32; do for wa in 0 1; do
  NGP_LANES=$lanes NGP_WALK_AHEAD=$wa timeout 600 python tools/bench_fox.py --steps 5000 --profile 32 --eval-views 2 > gpurun_out/s_fox_l${lanes}_wa${wa}.json 2> gpurun_out/s_fox_l${lanes}_wa${wa}.err
done; done
( timeout 900 python -m pytest tests/test_nerf_gpu.py -m gpu -x -q -k "walk_ahead or cascades" 2>&1 | tail -30 ) > gpurun_out/s_pytest_nerf.txt
echo done
