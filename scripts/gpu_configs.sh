#!/bin/bash
# BASELINE.json configs[3] (512^3, strong scaling over z-slabs) and configs[4] (1024x512x512, acc=200, 8 GPUs).
# bash scripts/gpu_configs.sh <tag> "<list of N for config 3>" [run config 4: 0/1]
TAG=${1:-cfg}; NS=${2:-"1 2 4 8"}; C4=${3:-1}
OUT=gpurun_out/$TAG; mkdir -p $OUT
run() { # n, name, args...
  n=$1; name=$2; shift 2
  if [ "$n" = 1 ]; then CMD="python bench.py"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py"; fi
  timeout 400 $CMD --gpus $n "$@" --no-cpu > $OUT/$name.json 2> $OUT/$name.err; echo "$name rc=$?"; cut -c1-200 $OUT/$name.json
}
for n in $NS; do run $n c3_n$n --size 512x512x512 --strong --steps 3 --warmup 2; done
if [ "$C4" = 1 ]; then run 8 c4_n8 --size 1024x512x512 --strong --acc 200 --steps 2 --warmup 1; fi
