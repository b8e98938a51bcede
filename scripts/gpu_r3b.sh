#!/bin/bash
# round 2, visit 3b (4 GPUs): thin slabs -- whole-column sweeps with the hand-over after the sweep (no face segments) against the default
mkdir -p gpurun_out/r3b
run() {  # tag, size args, env...
  tag=$1; shift; size=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 4 --steps 3 --warmup 2 --no-extra $size > gpurun_out/r3b/b_$tag.json 2> gpurun_out/r3b/b_$tag.err
  echo "rc=$?" >> gpurun_out/r3b/b_$tag.err
}
run thin_nooverlap "--size 512x512x256 --strong" FLUID_SW_OVERLAP=0
run weak_nooverlap "" FLUID_SW_OVERLAP=0
