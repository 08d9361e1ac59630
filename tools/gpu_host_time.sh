#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
( timeout 300 python tools/host_time.py; NGP_DEBUG_SKIP_MASK_SCAN=1 timeout 300 python tools/host_time.py ) > gpurun_out/host_time.txt 2>&1
cat gpurun_out/host_time.txt
