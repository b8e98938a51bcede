#!/bin/bash
# programmatic dependent launch between the sweeps of a solve: parity with it on, then bench off / on / off / on
TAG=${1:-ab4}; OUT=gpurun_out/$TAG; mkdir -p $OUT
FLUID_SW_PDL=1 timeout 200 python -m pytest tests/test_gpu_tiled.py tests/test_gpu_parity.py -m gpu -x -q > $OUT/pytest_pdl.log 2>&1; echo "pytest(pdl) rc=$?"; tail -2 $OUT/pytest_pdl.log
for v in 0 1 0 1; do
  FLUID_SW_PDL=$v timeout 100 python bench.py --steps 10 --warmup 3 --no-cpu > $OUT/bench_pdl$v.json 2> $OUT/bench.err
  python -c "
import json; d=json.load(open('$OUT/bench_pdl$v.json')); print('PDL=$v ms/step', round(d['ms_per_step'],3), 'us/sweep', round(d['roofline']['us_per_field_sweep'],2), 'e2e', round(d['e2e']['ms_per_step'],2), 'dens_sum', d['stats']['dens_sum'])"
done
