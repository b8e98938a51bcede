#!/bin/bash
# round 2, visit m (1 GPU, final code): full parity, bench with cpu baseline + extras, reference arm, ncu launch list + full capture
mkdir -p gpurun_out/r2m
python -m pytest tests -x -q -m gpu > gpurun_out/r2m/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2m/pytest.log
python bench.py > gpurun_out/r2m/bench.json 2> gpurun_out/r2m/bench.err
echo "rc=$?" >> gpurun_out/r2m/bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2m/ref.json 2> gpurun_out/r2m/ref.err
python scripts/kernels_bench.py > gpurun_out/r2m/kb.json 2> gpurun_out/r2m/kb.err
python scripts/sweep_bench.py --tag final > gpurun_out/r2m/sb.json 2>> gpurun_out/r2m/kb.err
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra"
$BENCH > gpurun_out/r2m/plain.json 2> gpurun_out/r2m/plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 790 -c 270 --csv --log-file gpurun_out/r2m/launches.csv $BENCH > gpurun_out/r2m/ncu_ll.log 2>&1
echo "ll rc=$?" >> gpurun_out/r2m/ncu_ll.log
CMD="python scripts/sweep_bench.py --acc 10 --reps 1 --tag ncu"
$CMD > gpurun_out/r2m/plain_sb.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_sweep_stream -s 12 -c 2 -o gpurun_out/r2m/sweep_zbits $CMD > gpurun_out/r2m/ncu_full.log 2>&1
echo "full rc=$?" >> gpurun_out/r2m/ncu_full.log
