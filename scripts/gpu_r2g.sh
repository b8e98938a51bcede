#!/bin/bash
# round 2, visit g (2 GPUs): slab parity with the 4-plane halos, bench at N=2 with slab_parity + extras
mkdir -p gpurun_out/r2g
python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2g/pytest_slab.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2g/pytest_slab.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2g/bench2.json 2> gpurun_out/r2g/bench2.err
echo "rc=$?" >> gpurun_out/r2g/bench2.err
