#!/bin/bash
# round 2, visit h (2 GPUs): push-based exchanges + flag barrier + 4-field slab diffusion: parity and A/B
mkdir -p gpurun_out/r2h
python -m pytest tests/test_gpu_slab.py -x -q -m gpu > gpurun_out/r2h/pytest_slab.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2h/pytest_slab.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 --no-extra > gpurun_out/r2h/bench2.json 2> gpurun_out/r2h/bench2.err
echo "rc=$?" >> gpurun_out/r2h/bench2.err
FLUID_XCHG_NCCL=1 $TR --master-port 29534 bench.py --gpus 2 --steps 5 --warmup 3 --no-extra > gpurun_out/r2h/bench2_nccl.json 2> gpurun_out/r2h/bench2_nccl.err
FLUID_NO_BATCH4=1 $TR --master-port 29535 bench.py --gpus 2 --steps 5 --warmup 3 --no-extra > gpurun_out/r2h/bench2_nobatch4.json 2> gpurun_out/r2h/bench2_nobatch4.err
