#!/bin/bash
mkdir -p gpurun_out
timeout 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err
tail -c 300 gpurun_out/z_bench.json
