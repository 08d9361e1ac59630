#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_render_frame_gpu.py tests/test_image_gpu.py -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/m_pytest_frame.txt
( timeout 1500 python -m pytest tests/test_nerf_gpu.py tests/test_snapshot_gpu.py tests/test_nerf_loader_gpu.py tests/test_psnr_parity_gpu.py -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/m_pytest_nerf.txt
timeout 900 python tools/bench_render.py 19 > gpurun_out/m_render_t19.json 2> gpurun_out/m_render_t19.err
NGP_TRACE_NO_COARSE=1 timeout 900 python tools/bench_render.py 19 > gpurun_out/m_render_t19_nocoarse.json 2> gpurun_out/m_render_t19_nocoarse.err
timeout 900 python tools/bench_fox.py --steps 3000 --profile 64 --eval-views 4 > gpurun_out/m_fox_3000.json 2> gpurun_out/m_fox_3000.err
echo done
