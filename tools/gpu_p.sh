#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_nerf_gpu.py tests/test_render_frame_gpu.py tests/test_snapshot_gpu.py -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/p_pytest.txt
for t in 2097152 8388608 16777216; do
  NGP_TRACE_TARGET_SAMPLES=$t timeout 600 python tools/render_timeline.py 19 > gpurun_out/p_render_timeline_$t.txt 2>&1
done
bash tools/gpu_fox_ab.sh
echo done
