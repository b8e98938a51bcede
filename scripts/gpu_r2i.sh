#!/bin/bash
# round 2, visit i (2 GPUs): slab parity with push exchanges, bisecting switches
mkdir -p gpurun_out/r2i
python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2i/pytest_slab.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2i/pytest_slab.log
FLUID_SLAB_LOG_TAG=_nccl FLUID_XCHG_NCCL=1 python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2i/pytest_slab_nccl.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2i/pytest_slab_nccl.log
FLUID_SLAB_LOG_TAG=_nobatch4 FLUID_NO_BATCH4=1 python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2i/pytest_slab_nobatch4.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2i/pytest_slab_nobatch4.log
