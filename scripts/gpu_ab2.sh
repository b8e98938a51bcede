#!/bin/bash
# A/B of the gradient (multiply by 1/h when h is a power of two) and advect (lanes along x) kernels, then the GPU suite.
TAG=${1:-ab2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 300 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
python scripts/kernels_bench.py --reps 20 > $OUT/kb_new.txt 2>&1; grep -E "gradient|advect" $OUT/kb_new.txt | head -4
FLUID_ADVECT_QUAD=1 FLUID_GRAD_DIV=1 python scripts/kernels_bench.py --reps 20 > $OUT/kb_old.txt 2>&1; grep -E "gradient|advect" $OUT/kb_old.txt | head -4
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('$OUT/bench.json')); print('ms/step', d['ms_per_step'], 'GLUPS', d['value'], 'e2e', d['e2e']['ms_per_step'], 'roof', d['roofline']['frac'])"
FLUID_ADVECT_QUAD=1 FLUID_GRAD_DIV=1 timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu > $OUT/bench_old.json 2> $OUT/bench_old.err; python -c "
import json; d=json.load(open('$OUT/bench_old.json')); print('OLD ms/step', d['ms_per_step'], 'GLUPS', d['value'], 'e2e', d['e2e']['ms_per_step'])"
