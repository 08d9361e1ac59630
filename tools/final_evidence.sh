set -x
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
NGP_NCU=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python bench.py --steps 4 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
NGP_NCU=1 ncu --profile-from-start off --set full --clock-control none --import-source on -f -o gpurun_out/prof_r1_final python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_full.log 2>&1
python tools/microbench.py > gpurun_out/microbench.json 2> gpurun_out/microbench.err
python tools/bench_configs.py > gpurun_out/configs.log 2>&1
python tools/bench_render.py 19 > gpurun_out/render.json 2> gpurun_out/render.err
ls -la gpurun_out | head -30
