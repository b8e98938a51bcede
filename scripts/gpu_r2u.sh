#!/bin/bash
# round 2, visit u (4 GPUs): slab parity with the buffer swap and the early face signal (separate edge CTAs and the one-wave plan),
# the C++ multi-GPU driver, and the two plans timed against each other
mkdir -p gpurun_out/r2u
python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2u/pytest_slab.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2u/pytest_slab.log
FLUID_SW_FACE_CTA=2 FLUID_SLAB_LOG_TAG=_onewave python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2u/pytest_slab_onewave.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2u/pytest_slab_onewave.log
python -m pytest tests/test_host_cpp.py -x -q -m gpu > gpurun_out/r2u/pytest_host.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2u/pytest_host.log
run() {  # tag, env...
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29591 bench.py --gpus 4 --steps 5 --warmup 3 --no-extra > gpurun_out/r2u/bench4_$tag.json 2> gpurun_out/r2u/bench4_$tag.err
  echo "rc=$?" >> gpurun_out/r2u/bench4_$tag.err
}
run default FLUID_X=0
run onewave FLUID_SW_FACE_CTA=2
run default_b FLUID_X=0
run onewave_b FLUID_SW_FACE_CTA=2
