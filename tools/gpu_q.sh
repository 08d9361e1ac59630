#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/q_pytest_all.txt
timeout 900 python tools/bench_render.py 19 > gpurun_out/q_render_t19.json 2> gpurun_out/q_render_t19.err
timeout 600 python bench.py --render 3840x2160 --steps 10 --warmup 3 > gpurun_out/q_bench_render4k.json 2> gpurun_out/q_bench_render4k.err
echo done
