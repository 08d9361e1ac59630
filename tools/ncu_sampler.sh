#!/bin/bash
# ncu --set full of the training marcher kernels at steady state, for lanes_per_ray variants (tools; run under gpurun)
set -e
python tools/phase_profile.py 100 800 1100 > gpurun_out/plain_sampler.log 2>&1
for L in "$@"; do
  NGP_LANES=$L NGP_NCU=1 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"sampler_" -c 2 -f -o gpurun_out/prof_sampler_L$L python tools/phase_profile.py 100 800 1100 > gpurun_out/ncu_sampler_L$L.log 2>&1
done
