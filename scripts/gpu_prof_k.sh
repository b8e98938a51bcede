#!/bin/bash
# ncu --set full of the non-sweep kernels through scripts/kernels_bench.py.  bash scripts/gpu_prof_k.sh <tag> <regex> [count]
TAG=$1; KRE=$2; CNT=${3:-2}
OUT=gpurun_out/$TAG; mkdir -p $OUT
CMD="python scripts/kernels_bench.py --reps 2"
$CMD > $OUT/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$KRE -s 4 -c $CNT -o $OUT/prof $CMD > $OUT/ncu.log 2>&1
echo "rc=$?"; tail -3 $OUT/ncu.log
