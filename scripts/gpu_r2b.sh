#!/bin/bash
# round 2, visit b: pair-sweep kernel parity + A/B timing
mkdir -p gpurun_out/r2b
python -m pytest tests/test_gpu_tiled.py tests/test_gpu_parity.py tests/test_gpu_golden.py -x -q -m gpu > gpurun_out/r2b/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b/pytest.log
python scripts/sweep_bench.py --tag pair > gpurun_out/r2b/sb_pair.json 2> gpurun_out/r2b/sb_pair.err
FLUID_NO_PAIR=1 python scripts/sweep_bench.py --tag nopair > gpurun_out/r2b/sb_nopair.json 2>> gpurun_out/r2b/sb_pair.err
FLUID_SW_PLAN=1 python scripts/sweep_bench.py --tag pair_uniform > gpurun_out/r2b/sb_pair_uniform.json 2>> gpurun_out/r2b/sb_pair.err
FLUID_SW_PDL=0 python scripts/sweep_bench.py --tag pair_nopdl > gpurun_out/r2b/sb_pair_nopdl.json 2>> gpurun_out/r2b/sb_pair.err
FLUID_SW_ZIGZAG=0 python scripts/sweep_bench.py --tag pair_nozz > gpurun_out/r2b/sb_pair_nozz.json 2>> gpurun_out/r2b/sb_pair.err
python bench.py --no-cpu > gpurun_out/r2b/bench.json 2> gpurun_out/r2b/bench.err
