#!/bin/bash
# round-2 GPU call C: TS-form forward MLP — parity of every path that runs it, then A/B against the SS form
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_image_gpu.py tests/test_sdf_gpu.py -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/c_pytest_ops.txt
( timeout 600 python -m pytest tests/test_nerf_gpu.py -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/c_pytest_nerf.txt
for ts in 0 1; do
  NGP_MLP_TS=$ts timeout 600 python bench.py --steps 64 --warmup 16 --repeats 7 --no-cpu-baseline > gpurun_out/c_bench_ts$ts.json 2> gpurun_out/c_bench_ts$ts.err
done
echo done
