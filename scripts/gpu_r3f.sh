#!/bin/bash
# round 2, visit 3f (2 GPUs): the slab swap as a one-time agreement between the ranks -- slab parity with and without it, C++ driver
mkdir -p gpurun_out/r3f
python -m pytest tests/test_gpu_slab.py tests/test_host_cpp.py -x -q -m gpu > gpurun_out/r3f/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3f/pytest.log
FLUID_NO_SLAB_SWAP=1 python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r3f/pytest_noswap.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3f/pytest_noswap.log
