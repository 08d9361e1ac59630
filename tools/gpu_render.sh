#!/bin/bash
# round 2: inference tracer timing, 1 GPU: 1080p / 4K, full frame + one band, per-kernel profile (tools/bench_render.py)
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
timeout 900 python tools/bench_render.py 19 > gpurun_out/l_render_t19.json 2> gpurun_out/l_render_t19.err
NGP_TRACE_NO_COARSE=1 timeout 900 python tools/bench_render.py 19 > gpurun_out/l_render_t19_nocoarse.json 2> gpurun_out/l_render_t19_nocoarse.err
timeout 600 python bench.py --render 3840x2160 --steps 10 --warmup 3 > gpurun_out/l_bench_render4k.json 2> gpurun_out/l_bench_render4k.err
echo done
