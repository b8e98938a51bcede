#!/bin/bash
# launch list of one bench step under ncu (gpu__time_duration only), after the same command ran plain
TAG=${1:-ll}; OUT=gpurun_out/$TAG; mkdir -p $OUT
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu"
timeout 40 $BENCH > $OUT/plain.json 2> $OUT/plain.err && \
timeout 50 ncu --metrics gpu__time_duration.sum --clock-control none -s 790 -c 270 --csv --log-file $OUT/launches.csv $BENCH > $OUT/ncu.log 2>&1
echo "rc=$?"; wc -l $OUT/launches.csv; tail -2 $OUT/ncu.log
