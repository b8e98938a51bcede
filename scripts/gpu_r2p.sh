#!/bin/bash
# round 2, visit p (1 GPU): marching cubes, gradient slow path out of line, ncu captures of the new kernels
mkdir -p gpurun_out/r2p
python -m pytest tests -x -q -m gpu > gpurun_out/r2p/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2p/pytest.log
python scripts/kernels_bench.py > gpurun_out/r2p/kb.json 2> gpurun_out/r2p/kb.err
python scripts/kernels_bench.py --d 512 > gpurun_out/r2p/kb_d512.json 2>> gpurun_out/r2p/kb.err
FLUID_ADVECT_NY=1 python scripts/kernels_bench.py > gpurun_out/r2p/kb_ny1.json 2>> gpurun_out/r2p/kb.err
CMD="python scripts/kernels_bench.py --d 512 --reps 2"
$CMD > gpurun_out/r2p/plain_kb.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_advect3|k_stats_quads|k_gradient' -s 9 -c 9 -o gpurun_out/r2p/kernels $CMD > gpurun_out/r2p/ncu_full.log 2>&1
echo "full rc=$?" >> gpurun_out/r2p/ncu_full.log
