# wall ms/step of the training step for marcher variants (lanes per ray G x candidates per lane K), ray prefetch on
for cfg in "8 1" "4 2" "4 1" "2 4" "2 8" "1 1" "32 1"; do set -- $cfg; echo "G=$1 K=$2"; NGP_NSTEP=96 NGP_LANES=$1 NGP_SAMPLER_K=$2 python tools/phase_profile.py 100 800 1100 2>&1 | grep step | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['ms_per_step_wall'], d['kernels_ms_per_step']['sampler'], d['measured'])"; done
