#!/bin/bash
# round 2, visit o (1 GPU): fused velocity advection, single-launch statistics, division-free gradient for non-power-of-two h
mkdir -p gpurun_out/r2o
python -m pytest tests -x -q -m gpu > gpurun_out/r2o/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2o/pytest.log
python scripts/kernels_bench.py > gpurun_out/r2o/kb.json 2> gpurun_out/r2o/kb.err
python scripts/kernels_bench.py --d 512 > gpurun_out/r2o/kb_d512.json 2>> gpurun_out/r2o/kb.err
FLUID_GRAD_DIV=1 FLUID_ADVECT_FUSED=0 python scripts/kernels_bench.py --d 512 > gpurun_out/r2o/kb_d512_old.json 2>> gpurun_out/r2o/kb.err
python bench.py --no-cpu --no-extra > gpurun_out/r2o/bench.json 2> gpurun_out/r2o/bench.err
FLUID_ADVECT_FUSED=0 python bench.py --no-cpu --no-extra > gpurun_out/r2o/bench_unfused.json 2>> gpurun_out/r2o/bench.err
