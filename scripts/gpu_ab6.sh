#!/bin/bash
# density diffusion batched as the fourth field of the velocity diffusion: GPU suite with it (default), bench with / without
TAG=${1:-ab6}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
for v in batch4 nobatch4; do
  if [ $v = nobatch4 ]; then export FLUID_NO_BATCH4=1; fi
  timeout 100 python bench.py --steps 10 --warmup 3 --no-cpu > $OUT/bench_$v.json 2> $OUT/bench.err
  python -c "
import json; d=json.load(open('$OUT/bench_$v.json')); print('$v ms/step', round(d['ms_per_step'],3), 'GLUPS', round(d['value'],4), 'us/sweep', round(d['roofline']['us_per_field_sweep'],2), 'frac', round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['ms_per_step'],2), 'launches', d['gpu_launches'], 'dens_sum', d['stats']['dens_sum'], 'vx_max', d['stats']['vx_max'])"
done
