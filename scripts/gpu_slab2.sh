#!/bin/bash
# 2-GPU visit: slab parity (new advect kernel in slab mode), the same with dependent launches inside slabs, bench A/B
TAG=${1:-slab2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_slab.py -m gpu -x -q > $OUT/pytest_slab.log 2>&1; echo "slab pytest rc=$?"; tail -2 $OUT/pytest_slab.log
FLUID_SW_PDL=2 timeout 300 python -m pytest tests/test_gpu_slab.py -m gpu -x -q > $OUT/pytest_slab_pdl.log 2>&1; echo "slab pytest (pdl) rc=$?"; tail -2 $OUT/pytest_slab_pdl.log
for v in 1 2 1 2; do
  FLUID_SW_PDL=$v timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu > $OUT/bench2_pdl$v.json 2> $OUT/bench2.err
  python -c "
import json; d=json.load(open('$OUT/bench2_pdl$v.json')); print('PDL=$v ms/step', round(d['ms_per_step'],3), 'GLUPS', round(d['value'],4), 'us/sweep', round(d['roofline']['us_per_field_sweep'],2), 'e2e', round(d['e2e']['ms_per_step'],2))"
done
