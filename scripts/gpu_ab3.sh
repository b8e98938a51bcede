#!/bin/bash
# advect variants: rows per thread 1 / 2 / 4 against the quad form; parity tests for the default first.
TAG=${1:-ab3}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tiled.py -m gpu -x -q -k "advect or step or run or slab" > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/pytest.log
for ny in 1 2 4; do echo "NY=$ny"; FLUID_ADVECT_NY=$ny python scripts/kernels_bench.py --reps 20 2>&1 | grep -E "advect" | head -3; done | tee $OUT/kb_ny.txt
echo QUAD; FLUID_ADVECT_QUAD=1 python scripts/kernels_bench.py --reps 20 2>&1 | grep -E "advect" | head -3 | tee $OUT/kb_quad.txt
