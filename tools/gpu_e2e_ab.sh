#!/bin/bash
# the upload path (e2e) of bench.py: host half of the step's end deferred (default) or inside train() (NGP_DEFER_FINISH=0)
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; : > gpurun_out/e2e_ab.txt
for cfg in "NGP_DEFER_FINISH=0" "NGP_DEFER_FINISH=1"; do
  echo "== $cfg" >> gpurun_out/e2e_ab.txt
  ( env $cfg timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value', round(b['value']/1e6,1), 'ms', round(b['ms_per_step'],4), 'e2e', round(b['e2e']['value']/1e6,1), 'e2e ms', round(b['e2e']['ms_per_step'],4), 'launches', b['gpu_launches'])
" ) >> gpurun_out/e2e_ab.txt
done
( timeout 300 python tools/host_time.py 2>&1 | tail -8 ) >> gpurun_out/e2e_ab.txt
( timeout 600 python -m pytest tests/test_nerf_gpu.py tests/test_exposure_gpu.py tests/test_snapshot_gpu.py -x -q -m gpu 2>&1 | tail -3 ) >> gpurun_out/e2e_ab.txt
cat gpurun_out/e2e_ab.txt
