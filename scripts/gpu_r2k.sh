#!/bin/bash
# round 2, visit k (8 GPUs): slab parity on 8 ranks (interior ranks have two faces), bench at N = 8 and N = 4 with the extras
mkdir -p gpurun_out/r2k
python -m pytest tests/test_gpu_slab.py -x -q -m gpu -k "8" > gpurun_out/r2k/pytest_slab8.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2k/pytest_slab8.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29541 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2k/bench8.json 2> gpurun_out/r2k/bench8.err
echo "rc=$?" >> gpurun_out/r2k/bench8.err
$TR --nproc-per-node 4 --master-port 29542 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2k/bench4.json 2> gpurun_out/r2k/bench4.err
echo "rc=$?" >> gpurun_out/r2k/bench4.err
FLUID_XCHG_NCCL=1 $TR --nproc-per-node 8 --master-port 29543 bench.py --gpus 8 --steps 5 --warmup 3 --no-extra > gpurun_out/r2k/bench8_nccl.json 2> gpurun_out/r2k/bench8_nccl.err
