#!/bin/bash
# 8-GPU runs of the round-2 bench (gpurun --gpus 8 -- bash tools/scale_r2.sh): the default line (sample split, tile
# split beside it, film check, C4 sub-record, single-process mode) and the tile split through the megakernel.
N=${N:-8}
TAG=${TAG:-r02i}
mkdir -p gpurun_out
run() { # name, extra args
  local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
      bench.py --gpus $N --steps 5 --warmup 3 "$@" > gpurun_out/${TAG}_${name}_n$N.json 2> gpurun_out/${TAG}_${name}_n$N.err
  echo "$name rc=$?"; cut -c1-160 gpurun_out/${TAG}_${name}_n$N.json
}
run bench_c3
run bench_c3_tile_mega --split tile --mode megakernel --no-extras
