#!/bin/bash
# round 2, visit c: ncu full capture of the pair-sweep kernel (single-field pressure solve, 256^3)
mkdir -p gpurun_out/r2c
CMD="python scripts/sweep_bench.py --acc 4 --reps 1 --tag ncu"
$CMD > gpurun_out/r2c/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_sweep_pair -s 2 -c 2 -o gpurun_out/r2c/pair_v1 $CMD > gpurun_out/r2c/ncu.log 2>&1
echo "rc=$?" >> gpurun_out/r2c/ncu.log
