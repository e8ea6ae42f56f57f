#!/bin/bash
# round-2 GPU call 2 (2 GPUs): full GPU regression incl. the 2-rank == 1-rank tests, bench smoke at 128^3, N=2 bench
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c2_gpus.log 2>&1
timeout 300 python bench.py --cells 128 --steps 3 --warmup 3 --e2e-long 5 --cpu-cells 32 > gpurun_out/c2_bench128.json 2> gpurun_out/c2_bench128.err
echo "rc=$?" >> gpurun_out/c2_bench128.err
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 2>&1 | tail -15 > gpurun_out/c2_tests.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --cells 128 --steps 5 --warmup 3 --no-cpu > gpurun_out/c2_bench128_n2.json 2> gpurun_out/c2_bench128_n2.err
echo "rc=$?" >> gpurun_out/c2_bench128_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29702 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu > gpurun_out/c2_bench512_n2.json 2> gpurun_out/c2_bench512_n2.err
echo "rc=$?" >> gpurun_out/c2_bench512_n2.err
ls -la gpurun_out | tail -12
