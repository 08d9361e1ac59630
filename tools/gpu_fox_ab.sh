#!/bin/bash
# config 3 (fox, aabb_scale 8: cascades + cone angle 1/256), 35 000 steps: the product library against the A/B build with the marcher's
# logf / expf in Horner form and IEEE divisions (csrc/Makefile target `ab`)
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
timeout 900 python tools/bench_fox.py --steps 35000 --profile 64 --eval-views 10 > gpurun_out/o_fox_35k_new.json 2> gpurun_out/o_fox_35k_new.err
NGP_B200_LIB=$PWD/instant-ngp_b200/libngp_b200_ab.so timeout 900 python tools/bench_fox.py --steps 35000 --profile 64 --eval-views 10 > gpurun_out/o_fox_35k_oldmath.json 2> gpurun_out/o_fox_35k_oldmath.err
echo done
