#!/bin/bash
# round 2, visit 3e (2 GPUs, final code): slab parity, C++ multi-GPU driver (frames against the single-GPU run), python frame files
mkdir -p gpurun_out/r3e
python -m pytest tests/test_gpu_slab.py tests/test_host_cpp.py -x -q -m gpu > gpurun_out/r3e/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3e/pytest.log
