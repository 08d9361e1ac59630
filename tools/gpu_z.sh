#!/bin/bash
# scratch wrapper for one gpurun call: edit the selection, run `bash tools/gpurun_retry.sh gpurun_out/z_call.log 1800 'bash tools/gpu_z.sh'`
mkdir -p gpurun_out
timeout 100 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/z_smoke.txt 2>&1
tail -2 gpurun_out/z_smoke.txt
