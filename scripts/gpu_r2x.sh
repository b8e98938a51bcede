#!/bin/bash
# round 2, visit x (1 GPU): row tiles of 16 owned rows (<2,9>: 576 threads, 16 tiles at H = 256 instead of 19) against <2,8>
mkdir -p gpurun_out/r2x
FLUID_SW_CFG=4 python -m pytest tests/test_gpu_tiled.py -x -q -m gpu > gpurun_out/r2x/pytest_cfg4.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2x/pytest_cfg4.log
python bench.py --no-cpu --no-extra > gpurun_out/r2x/bench_default.json 2> gpurun_out/r2x/bench.err
FLUID_SW_CFG=4 python bench.py --no-cpu --no-extra > gpurun_out/r2x/bench_cfg4.json 2>> gpurun_out/r2x/bench.err
FLUID_SW_CFG=4 python scripts/sweep_bench.py --tag cfg4 > gpurun_out/r2x/sb_cfg4.json 2>> gpurun_out/r2x/bench.err
python scripts/sweep_bench.py --tag default > gpurun_out/r2x/sb_default.json 2>> gpurun_out/r2x/bench.err
