#!/bin/bash
# A/B of the slab sweep variants at N GPUs.  bash scripts/gpu_slab_ab.sh <tag> <N> <size> "<env1>" "<env2>" ...
TAG=$1; N=$2; SIZE=$3; shift 3
OUT=gpurun_out/$TAG; mkdir -p $OUT
i=0
for cfg in "$@"; do
  i=$((i+1))
  env $cfg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --size $SIZE --strong --steps 3 --warmup 3 --no-cpu > $OUT/r$i.json 2> $OUT/r$i.err
  echo "== $cfg rc=$?"
  python - <<PY
import json
d=json.loads(open("$OUT/r$i.json").read().strip().splitlines()[-1])
print("ms/step %.1f  us/field-sweep %.1f  solver share %.3f"%(d["ms_per_step"], d["roofline"]["us_per_field_sweep"], d["roofline"]["solver_share_of_step"]))
PY
done
