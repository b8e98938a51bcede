#!/bin/bash
# round 2, visit 3g (2 GPUs): the plan builder moved into sweep_plan.h (device code byte-identical) -- streaming-kernel parity,
# slab parity, one bench line on two slabs
mkdir -p gpurun_out/r3g
python -m pytest tests/test_gpu_tiled.py tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r3g/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3g/pytest.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 2 --steps 3 --warmup 3 --no-extra > gpurun_out/r3g/bench2.json 2> gpurun_out/r3g/bench2.err
echo "rc=$?" >> gpurun_out/r3g/bench2.err
