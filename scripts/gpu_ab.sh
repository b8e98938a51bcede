#!/bin/bash
# tiled tests + sweep micro-benchmark under a few env settings.  bash scripts/gpu_ab.sh <tag> "<env1>" "<env2>" ...
TAG=${1:-ab}; shift; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_tiled.py -x -q > $OUT/pytest_tiled.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_tiled.log
for cfg in "$@"; do
  echo "== $cfg"; env $cfg python scripts/sweep_bench.py --acc 40 | tee -a $OUT/ab.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('single %.1f us (%.2f)  mix %.1f us (%.2f)'%(d['single']['us_per_sweep'],d['single']['frac'],d['step_mix']['us_per_field_sweep'],d['step_mix']['frac']))"
done
