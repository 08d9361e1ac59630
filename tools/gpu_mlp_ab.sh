#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; : > gpurun_out/mlp_ab.txt
( timeout 120 python -m pytest tests/test_ops_gpu.py tests/test_image_gpu.py tests/test_sdf_gpu.py -m gpu -x -q -k "forward_backward or image_network or image_training or sdf or inference" 2>&1 | tail -4 ) >> gpurun_out/mlp_ab.txt
for cfg in "NGP_X=0" $EXTRA_CFGS; do
  ( env $cfg timeout 120 python tools/bench_mlp.py 2>&1 | tail -2 ) >> gpurun_out/mlp_ab.txt
done
( NGP_MLP_PROBE=1 REPS=3 timeout 120 python tools/bench_mlp.py 2>&1 | grep "backward probe" | tail -3 ) >> gpurun_out/mlp_ab.txt
