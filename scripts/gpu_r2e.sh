#!/bin/bash
# round 2, visit e: deeper halos (layout change) + pair-sweep v4 (branch-free stages): full parity, timing, ncu
mkdir -p gpurun_out/r2e
python -m pytest tests -x -q -m gpu > gpurun_out/r2e/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2e/pytest.log
python scripts/sweep_bench.py --tag u6 > gpurun_out/r2e/sb_u6.json 2> gpurun_out/r2e/sb.err
FLUID_PAIR_UNROLL=2 python scripts/sweep_bench.py --tag u2 > gpurun_out/r2e/sb_u2.json 2>> gpurun_out/r2e/sb.err
FLUID_NO_PAIR=1 python scripts/sweep_bench.py --tag nopair > gpurun_out/r2e/sb_nopair.json 2>> gpurun_out/r2e/sb.err
python bench.py --no-cpu > gpurun_out/r2e/bench.json 2> gpurun_out/r2e/bench.err
CMD="python scripts/sweep_bench.py --acc 4 --reps 1 --tag ncu"
$CMD > gpurun_out/r2e/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_sweep_pair -s 2 -c 2 -o gpurun_out/r2e/pair_v4 $CMD > gpurun_out/r2e/ncu.log 2>&1
echo "rc=$?" >> gpurun_out/r2e/ncu.log
