#!/bin/bash
# round 2, visit n (1 GPU): the allocation-free statistics entry point -- alone in a fresh process, then after the whole GPU suite
mkdir -p gpurun_out/r2n
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2n/alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/r2n/alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/r2n/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2n/pytest.log
