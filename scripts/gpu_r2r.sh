#!/bin/bash
# round 2, visit r (4 GPUs): what a CTA that starts behind an edge CTA should be charged; edge depth; NUMA-local pinned buffers
mkdir -p gpurun_out/r2r
run() {  # tag, env...
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 4 --steps 5 --warmup 3 --no-extra > gpurun_out/r2r/bench4_$tag.json 2> gpurun_out/r2r/bench4_$tag.err
  echo "rc=$?" >> gpurun_out/r2r/bench4_$tag.err
}
run default FLUID_X=0
run nobind FLUID_NO_BIND=1
run late4 FLUID_SW_LATE=4
run late8 FLUID_SW_LATE=8
run late12 FLUID_SW_LATE=12
run edge4 FLUID_SW_EDGE=4
run edge4_late6 FLUID_SW_EDGE=4 FLUID_SW_LATE=6
nvidia-smi topo -m > gpurun_out/r2r/topo.txt 2>&1
