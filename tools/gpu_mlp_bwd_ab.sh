#!/bin/bash
# A/B of the fused MLP backward forms (NGP_MLP_BWD = 0 SS, 1 TS, 2 TS + activations kept in registers): parity tests under each, then timing
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; : > gpurun_out/mlp_bwd_ab.txt
for f in 0 1 2; do
  echo "== NGP_MLP_BWD=$f" >> gpurun_out/mlp_bwd_ab.txt
  ( NGP_MLP_BWD=$f timeout 200 python -m pytest tests/test_ops_gpu.py tests/test_image_gpu.py tests/test_sdf_gpu.py -m gpu -x -q -k "forward_backward or input_gradients or image_network or image_training or sdf" 2>&1 | tail -4 ) >> gpurun_out/mlp_bwd_ab.txt
  ( NGP_MLP_BWD=$f timeout 120 python tools/bench_mlp.py 2>&1 | tail -2 ) >> gpurun_out/mlp_bwd_ab.txt
  ( NGP_MLP_BWD=$f NGP_MLP_PROBE=1 REPS=3 timeout 120 python tools/bench_mlp.py 2>&1 | grep "backward probe" | tail -1 ) >> gpurun_out/mlp_bwd_ab.txt
done
cat gpurun_out/mlp_bwd_ab.txt
