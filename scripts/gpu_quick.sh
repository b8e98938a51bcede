#!/bin/bash
# quick GPU visit: tiled-kernel parity tests + one bench line.  bash scripts/gpu_quick.sh <tag> [pytest args]
TAG=${1:-q}; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_tiled.py -x -q "$@" > $OUT/pytest_tiled.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest_tiled.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; tail -3 $OUT/bench.err
python - <<PY
import json
d=json.load(open("$OUT/bench.json"))
print("ms/step", d["ms_per_step"], "GLUPS", d["value"], "e2e", d["e2e"]["value"], "roof", d["roofline"]["frac"], "us/sweep", d["roofline"]["us_per_field_sweep"], "solver share", d["roofline"]["solver_share_of_step"])
PY
