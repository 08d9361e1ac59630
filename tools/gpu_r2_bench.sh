#!/bin/bash
# round-2 evidence: both bench arms as the driver runs them (timed), ncu launch list + full capture of the same command
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
nproc > gpurun_out/i_nproc.txt
( time timeout 1500 python bench.py > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err ) 2> gpurun_out/i_bench.time
( time timeout 1500 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/i_ref.json 2> gpurun_out/i_ref.err ) 2> gpurun_out/i_ref.time
NGP_NCU=1 timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/i_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/i_ncu_launch.log 2>&1
NGP_NCU=1 timeout 1500 ncu --profile-from-start off --set full --clock-control none --import-source on -f -o gpurun_out/prof_r2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/i_ncu_full.log 2>&1
ls -la gpurun_out | head -30
