#!/bin/bash
# round 2, visit f: zero-bit volume in the single-sweep kernel: full parity + timing
mkdir -p gpurun_out/r2f
python -m pytest tests -x -q -m gpu > gpurun_out/r2f/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2f/pytest.log
python scripts/sweep_bench.py --tag zbits > gpurun_out/r2f/sb.json 2> gpurun_out/r2f/sb.err
python scripts/kernels_bench.py > gpurun_out/r2f/kb.json 2>> gpurun_out/r2f/sb.err
python bench.py --no-cpu > gpurun_out/r2f/bench.json 2> gpurun_out/r2f/bench.err
