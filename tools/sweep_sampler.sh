timeout 500 python -m pytest tests/test_nerf_gpu.py -m gpu -q --timeout 300 2>&1 | tail -2
for cfg in "8 1" "8 2" "4 2" "4 4" "4 1" "2 4" "2 8"; do set -- $cfg; echo "G=$1 K=$2"; NGP_LANES=$1 NGP_SAMPLER_K=$2 python tools/phase_profile.py 100 800 1100 2>&1 | grep step | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['ms_per_step_wall'], d['kernels_ms_per_step']['sampler'], d['measured'])"; done
