#!/bin/bash
# round 2, visit 3c (2 GPUs): the staged (pinned) obstacle upload of the e2e loop, on one GPU and on two slabs
mkdir -p gpurun_out/r3c
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "staged or advect_velocity" > gpurun_out/r3c/pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3c/pytest.log
python bench.py --no-cpu --no-extra > gpurun_out/r3c/bench1.json 2> gpurun_out/r3c/bench1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 2 --steps 5 --warmup 3 --no-extra > gpurun_out/r3c/bench2.json 2> gpurun_out/r3c/bench2.err
echo "rc=$?" >> gpurun_out/r3c/bench2.err
