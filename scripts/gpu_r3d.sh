#!/bin/bash
# round 2, visit 3d (1 GPU): frame snapshot kernel with eight loads in flight -- full parity, bench, launch list
mkdir -p gpurun_out/r3d
python -m pytest tests -x -q -m gpu > gpurun_out/r3d/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3d/pytest.log
python bench.py --no-cpu --no-extra > gpurun_out/r3d/bench.json 2> gpurun_out/r3d/bench.err
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra"
$BENCH > gpurun_out/r3d/plain.json 2> gpurun_out/r3d/plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 790 -c 270 --csv --log-file gpurun_out/r3d/launches.csv $BENCH > gpurun_out/r3d/ncu_ll.log 2>&1
echo "ll rc=$?" >> gpurun_out/r3d/ncu_ll.log
