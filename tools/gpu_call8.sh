#!/bin/bash
# round-2 GPU call 8 (8 GPUs): the strong-scaling line at N = 8 (global 512^3 cut into 8 z slabs), device-timed + e2e + parity check
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c8_gpus.log 2>&1
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 8 --steps 20 --warmup 3 --one-mode > gpurun_out/c13_bench512_n8.json 2> gpurun_out/c13_bench512_n8.err
echo "rc=$?" >> gpurun_out/c13_bench512_n8.err
tail -3 gpurun_out/c13_bench512_n8.err; tail -c 1500 gpurun_out/c13_bench512_n8.json
