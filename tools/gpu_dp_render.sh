#!/bin/bash
# round 2, after the tracer rewrite: tile-sharded 3840x2160 render on N GPUs, interleaved rows vs contiguous bands, T=2^19 and T=2^24; one training line per table size
N=${1:-8}; cd "$(dirname "$0")/.."; mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) bench.py --gpus $N "$@" > gpurun_out/dpr${N}_$tag.json 2> gpurun_out/dpr${N}_$tag.err; echo "$tag rc=$?"; }
run render4k_t19_interleaved --steps 9 --warmup 3 --render 3840x2160
run render4k_t19_bands --steps 9 --warmup 3 --render 3840x2160 --render-bands
run render4k_t24_interleaved --steps 9 --warmup 3 --render 3840x2160 --log2-hashmap-size 24
run t19_sharded --steps 20 --warmup 5 --repeats 7 --sharded-optimizer
run t24_sharded --steps 20 --warmup 5 --repeats 5 --log2-hashmap-size 24 --sharded-optimizer
echo done
