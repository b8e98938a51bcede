#!/bin/bash
# round 2, visit q (4 GPUs): slab sweeps -- early face signal, hand-over inside ordinary segments, dependent launch
mkdir -p gpurun_out/r2q
run() {  # tag, env...
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 4 --steps 5 --warmup 3 --no-extra > gpurun_out/r2q/bench4_$tag.json 2> gpurun_out/r2q/bench4_$tag.err
  echo "rc=$?" >> gpurun_out/r2q/bench4_$tag.err
}
run default FLUID_X=0
run latesignal FLUID_SW_DBG=16
run inseg FLUID_SW_FACE_CTA=0
run inseg_pdl FLUID_SW_FACE_CTA=0 FLUID_SW_PDL=2
run pdl FLUID_SW_PDL=2
