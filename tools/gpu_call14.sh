#!/bin/bash
# round-2 GPU call 14 (2 GPUs): the driver's N = 2 invocation shape (strong line + extra weak run) at a small size with the chunked-upload path forced
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29751 bench.py --gpus 2 --steps 5 --warmup 3 --cells 128 --host-gb 0.05 > gpurun_out/c14_bench128_n2.json 2> gpurun_out/c14_bench128_n2.err
echo "rc=$?" >> gpurun_out/c14_bench128_n2.err
tail -3 gpurun_out/c14_bench128_n2.err; tail -c 1200 gpurun_out/c14_bench128_n2.json
