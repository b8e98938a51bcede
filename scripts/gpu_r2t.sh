#!/bin/bash
# round 2, visit t (1 GPU): measured rebalancing of the sweep plan -- parity, then the bench with 0 / 3 / 6 rounds, with traces
mkdir -p gpurun_out/r2t
python -m pytest tests/test_gpu_tiled.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r2t/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2t/pytest.log
for r in 0 3 6; do
  FLUID_SW_TUNE=$r FLUID_SW_TRACE=gpurun_out/r2t/trace_tune$r FLUID_SW_TRACE_AT=1480,1560 python bench.py --steps 5 --warmup 3 --no-cpu --no-extra > gpurun_out/r2t/bench_tune$r.json 2> gpurun_out/r2t/bench_tune$r.err
done
FLUID_SW_TUNE=0 python scripts/sweep_bench.py --tag tune0 > gpurun_out/r2t/sb_tune0.json 2> gpurun_out/r2t/sb.err
FLUID_SW_TUNE=4 python scripts/sweep_bench.py --tag tune4 > gpurun_out/r2t/sb_tune4.json 2>> gpurun_out/r2t/sb.err
