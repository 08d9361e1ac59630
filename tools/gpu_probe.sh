#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
NGP_MLP_PROBE=1 timeout 300 python bench.py --steps 8 --warmup 4 --repeats 1 --pretrain 600 --no-cpu-baseline > gpurun_out/probe.json 2> gpurun_out/probe.err
grep "probe" gpurun_out/probe.err | tail -5 > gpurun_out/probe.txt
