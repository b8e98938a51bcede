#!/bin/bash
# ncu --set full capture of the sweep kernel through the micro-benchmark.  bash scripts/gpu_prof.sh <tag> <kernel-regex> [bench args]
TAG=$1; KRE=$2; shift 2
OUT=gpurun_out/$TAG; mkdir -p $OUT
CMD="python scripts/sweep_bench.py --acc 10 --reps 1 $@"
$CMD > $OUT/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$KRE -s 12 -c 2 -o $OUT/prof $CMD > $OUT/ncu.log 2>&1
echo "rc=$?"; cat $OUT/plain.log; tail -3 $OUT/ncu.log
