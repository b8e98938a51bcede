#!/bin/bash
# round 2, visit w (1 GPU, final code): full parity, bench with cpu baseline + extras, reference arm, per-kernel timings,
# ncu launch list + full capture of the sweep
mkdir -p gpurun_out/r2w
python -m pytest tests -x -q -m gpu > gpurun_out/r2w/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2w/pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2w/smoke.log 2>&1
python bench.py > gpurun_out/r2w/bench.json 2> gpurun_out/r2w/bench.err
echo "rc=$?" >> gpurun_out/r2w/bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2w/ref.json 2> gpurun_out/r2w/ref.err
python scripts/kernels_bench.py > gpurun_out/r2w/kb.json 2> gpurun_out/r2w/kb.err
python scripts/sweep_bench.py --tag final > gpurun_out/r2w/sb.json 2>> gpurun_out/r2w/kb.err
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra"
$BENCH > gpurun_out/r2w/plain.json 2> gpurun_out/r2w/plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 790 -c 270 --csv --log-file gpurun_out/r2w/launches.csv $BENCH > gpurun_out/r2w/ncu_ll.log 2>&1
echo "ll rc=$?" >> gpurun_out/r2w/ncu_ll.log
