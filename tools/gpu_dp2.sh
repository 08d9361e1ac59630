#!/bin/bash
# 2-GPU sanity of the data-parallel bench line as the driver launches it (+ the driver's own 1-GPU command)
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/dp2_bench.json 2> gpurun_out/dp2_bench.err
tail -c 600 gpurun_out/dp2_bench.json
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/n1_s20_bench.json 2> gpurun_out/n1_s20_bench.err
tail -c 600 gpurun_out/n1_s20_bench.json
