#!/bin/bash
# round 2, visit s (2 GPUs): per-CTA timelines of a slab sweep against the single-GPU sweep; buffer swap and the
# one-wave face plan in z-slabs (parity + timing)
mkdir -p gpurun_out/r2s
python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2s/pytest_slab.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2s/pytest_slab.log
FLUID_SW_FACE_CTA=2 FLUID_SLAB_LOG_TAG=_onewave python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2s/pytest_slab_onewave.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2s/pytest_slab_onewave.log
FLUID_SW_TRACE=gpurun_out/r2s/trace_n1 FLUID_SW_TRACE_AT=760,840 python bench.py --steps 3 --warmup 3 --no-cpu --no-extra > gpurun_out/r2s/bench1.json 2> gpurun_out/r2s/bench1.err
run() {  # tag, env...
  tag=$1; shift
  env "$@" FLUID_SW_TRACE=gpurun_out/r2s/trace_$tag FLUID_SW_TRACE_AT=760,840 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 2 --steps 3 --warmup 3 --no-extra > gpurun_out/r2s/bench2_$tag.json 2> gpurun_out/r2s/bench2_$tag.err
  echo "rc=$?" >> gpurun_out/r2s/bench2_$tag.err
}
run default FLUID_X=0
run noswap FLUID_NO_SLAB_SWAP=1
run onewave FLUID_SW_FACE_CTA=2
run onewave_late4 FLUID_SW_FACE_CTA=2 FLUID_SW_LATE=4
run late4 FLUID_SW_LATE=4
