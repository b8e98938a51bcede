#!/bin/bash
# One GPU-box visit: parity tests, bench line, then ONE ncu pass (a visit runs at most one profiler):
#   bash scripts/gpu_check.sh <tag> launches   -> launch list of the bench command (gpu__time_duration)
#   bash scripts/gpu_check.sh <tag> full       -> --set full capture of two k_sweep_stream launches
#   bash scripts/gpu_check.sh <tag> none       -> tests + bench only
TAG=${1:-r1}
PASS=${2:-launches}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
if [ "$PASS" != full ]; then
  python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
  tail -3 $OUT/pytest_gpu.log
  python bench.py --steps 5 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
  cat $OUT/bench.json
fi
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu"
if [ "$PASS" = launches ]; then
  $BENCH > $OUT/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 700 --csv \
      --log-file $OUT/launches.csv $BENCH > $OUT/ncu_launches.log 2>&1
  echo "ncu launches rc=$?"
elif [ "$PASS" = full ]; then
  $BENCH > $OUT/plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_sweep_stream -s 300 -c 2 \
      -o $OUT/prof_sweep $BENCH > $OUT/ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
ls -la $OUT
