#!/bin/bash
# round 2, visit l (2 GPUs): C++ multi-GPU driver against the single-GPU run; bench at N = 2 with the final code
mkdir -p gpurun_out/r2l
python -m pytest tests/test_host_cpp.py -x -q -m gpu > gpurun_out/r2l/pytest_host.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2l/pytest_host.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2l/bench2.json 2> gpurun_out/r2l/bench2.err
echo "rc=$?" >> gpurun_out/r2l/bench2.err
