#!/bin/bash
# round-2 GPU call 6 (2 GPUs): full GPU regression incl. the 2-rank == 1-rank tests and the thermal walls, then the driver's own N = 2 bench invocation
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c6_gpus.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q --durations=8 2>&1 | tail -25 > gpurun_out/c6_tests.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c6_bench512_n2.json 2> gpurun_out/c6_bench512_n2.err
echo "rc=$?" >> gpurun_out/c6_bench512_n2.err
tail -5 gpurun_out/c6_tests.log; tail -3 gpurun_out/c6_bench512_n2.err
