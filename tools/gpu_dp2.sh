#!/bin/bash
# 2-GPU sanity of the data-parallel bench line as the driver launches it
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/dp2_bench.json 2> gpurun_out/dp2_bench.err
tail -c 400 gpurun_out/dp2_bench.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/dp2_ref.json 2> gpurun_out/dp2_ref.err
tail -c 300 gpurun_out/dp2_ref.json
