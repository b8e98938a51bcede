#!/bin/bash
# Short GPU-box visit after the viewer kernels went in: their parity tests first, one bench line, then the whole
# GPU suite, each under its own timeout.   bash scripts/gpu_final.sh <tag>
TAG=${1:-final}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
t0=$(date +%s)
timeout 240 python -m pytest tests/test_postproc.py -m gpu -x -q > $OUT/pytest_postproc.log 2>&1; echo "postproc pytest rc=$? ($(( $(date +%s) - t0 )) s)"; tail -5 $OUT/pytest_postproc.log
timeout 120 python scripts/postproc_bench.py > $OUT/postproc_bench.json 2> $OUT/postproc_bench.err; echo "postproc bench rc=$?"; cat $OUT/postproc_bench.json
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$? ($(( $(date +%s) - t0 )) s)"; cat $OUT/bench.json
timeout 150 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
timeout ${2:-420} python -m pytest tests -m gpu -x -q --deselect tests/test_postproc.py > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$? ($(( $(date +%s) - t0 )) s)"; tail -4 $OUT/pytest_gpu.log
