#!/bin/bash
# round 2, visit z (8 GPUs, final code): weak-scaling line at N = 8 (slab parity inside), no extras
mkdir -p gpurun_out/r2z
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus 8 --steps 5 --warmup 3 --no-extra > gpurun_out/r2z/bench8.json 2> gpurun_out/r2z/bench8.err
echo "rc=$?" >> gpurun_out/r2z/bench8.err
nvidia-smi topo -m > gpurun_out/r2z/topo.txt 2>&1
