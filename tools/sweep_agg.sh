for a in 0 2 4 8 12; do
export NGP_AGG_MAX_RUNS=$a
python bench.py 2>/dev/null | python -c "
import sys,json,os
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('agg', os.environ['NGP_AGG_MAX_RUNS'], 'synthetic step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'grid_bwd', d['kernels']['grid_backward']['ms_per_step'])"
python tools/bench_fox.py --steps 1000 --spp 1 --eval-views 1 --profile 32 2>/dev/null | python -c "
import sys,json,os
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('agg', os.environ['NGP_AGG_MAX_RUNS'], 'fox on/off', d['ms_per_step_prefetch_on'], d['ms_per_step_prefetch_off'], 'grid_bwd', d['profile_prefetch_off']['kernels_ms_per_step']['grid_backward'])"
done
