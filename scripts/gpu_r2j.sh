#!/bin/bash
# round 2, visit j: x-tiled single-sweep kernel (W > 256): parity + timing at W = 512 and W = 1024
mkdir -p gpurun_out/r2j
python -m pytest tests -x -q -m gpu > gpurun_out/r2j/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2j/pytest.log
python scripts/sweep_bench.py --w 512 --h 512 --d 256 --acc 8 --reps 2 --tag w512_xtiles > gpurun_out/r2j/sb_512.json 2> gpurun_out/r2j/sb.err
FLUID_SW_CFG=4 python scripts/sweep_bench.py --w 512 --h 512 --d 256 --acc 8 --reps 2 --tag w512_cfg4 > gpurun_out/r2j/sb_512_cfg4.json 2>> gpurun_out/r2j/sb.err
python scripts/sweep_bench.py --w 1024 --h 512 --d 64 --acc 8 --reps 2 --tag w1024_xtiles > gpurun_out/r2j/sb_1024.json 2>> gpurun_out/r2j/sb.err
python scripts/sweep_bench.py --w 1024 --h 1024 --d 128 --acc 8 --reps 2 --tag w1024sq_xtiles > gpurun_out/r2j/sb_1024sq.json 2>> gpurun_out/r2j/sb.err
python bench.py --no-cpu > gpurun_out/r2j/bench.json 2> gpurun_out/r2j/bench.err
