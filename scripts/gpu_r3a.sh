#!/bin/bash
# round 2, visit 3a (4 GPUs): thin slabs -- 512 x 512 x 256 over 4 ranks = 64 planes per rank, the shape of configs[3] (512^3) on 8 GPUs:
# edge depth and where the face segments run
mkdir -p gpurun_out/r3a
run() {  # tag, env...
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 4 --steps 3 --warmup 2 --no-extra --size 512x512x256 --strong > gpurun_out/r3a/b_$tag.json 2> gpurun_out/r3a/b_$tag.err
  echo "rc=$?" >> gpurun_out/r3a/b_$tag.err
}
run default FLUID_X=0
run edge4 FLUID_SW_EDGE=4
run edge2 FLUID_SW_EDGE=2
run inseg FLUID_SW_FACE_CTA=0
run late0 FLUID_SW_LATE=0
