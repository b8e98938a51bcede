#!/bin/bash
# round 2, visit d: pair-sweep v3 (bank-conflict-free loads, static register ring): parity, timing A/B, ncu
mkdir -p gpurun_out/r2d
python -m pytest tests/test_gpu_tiled.py -x -q -m gpu -k "pair or headline" > gpurun_out/r2d/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2d/pytest.log
python scripts/sweep_bench.py --tag u6 > gpurun_out/r2d/sb_u6.json 2> gpurun_out/r2d/sb.err
FLUID_PAIR_UNROLL=2 python scripts/sweep_bench.py --tag u2 > gpurun_out/r2d/sb_u2.json 2>> gpurun_out/r2d/sb.err
FLUID_SW_PLAN=1 python scripts/sweep_bench.py --tag u6_uniform > gpurun_out/r2d/sb_u6_uniform.json 2>> gpurun_out/r2d/sb.err
FLUID_SW_PLAN=1 FLUID_PAIR_UNROLL=2 python scripts/sweep_bench.py --tag u2_uniform > gpurun_out/r2d/sb_u2_uniform.json 2>> gpurun_out/r2d/sb.err
python bench.py --no-cpu > gpurun_out/r2d/bench.json 2> gpurun_out/r2d/bench.err
CMD="python scripts/sweep_bench.py --acc 4 --reps 1 --tag ncu"
$CMD > gpurun_out/r2d/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_sweep_pair -s 2 -c 2 -o gpurun_out/r2d/pair_v3 $CMD > gpurun_out/r2d/ncu.log 2>&1
echo "rc=$?" >> gpurun_out/r2d/ncu.log
