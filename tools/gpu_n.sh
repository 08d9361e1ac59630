#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
timeout 600 python tools/render_timeline.py 19 > gpurun_out/n_render_timeline.txt 2>&1
( timeout 900 python -m pytest tests/test_render_frame_gpu.py -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/n_pytest_frame.txt
echo done
