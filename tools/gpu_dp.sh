#!/bin/bash
# data-parallel evidence on N GPUs of one box: dense all-reduce vs sharded optimizer at T=2^19 and T=2^24, tile-sharded 4K render
N=${1:-2}; cd "$(dirname "$0")/.."; mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) bench.py --gpus $N "$@" > gpurun_out/dp${N}_$tag.json 2> gpurun_out/dp${N}_$tag.err; echo "$tag rc=$?"; }
run t19_dense --steps 20 --warmup 5 --repeats 7 --dense-allreduce
run t19_sharded --steps 20 --warmup 5 --repeats 7 --sharded-optimizer
run t24_dense --steps 20 --warmup 5 --repeats 5 --log2-hashmap-size 24 --dense-allreduce
run t24_sharded --steps 20 --warmup 5 --repeats 5 --log2-hashmap-size 24 --sharded-optimizer
run render4k_t24 --steps 7 --warmup 2 --render 3840x2160 --log2-hashmap-size 24
run render4k_t19 --steps 7 --warmup 2 --render 3840x2160
if [ "$N" = "2" ]; then timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --repeats 5 --log2-hashmap-size 24 --no-cpu-baseline > gpurun_out/dp1_t24.json 2> gpurun_out/dp1_t24.err; timeout 300 python bench.py --steps 7 --warmup 2 --render 3840x2160 > gpurun_out/dp1_render4k_t19.json 2>gpurun_out/dp1_render.err; fi
nvidia-smi topo -m > gpurun_out/dp${N}_topo.txt 2>&1
echo done
