#!/bin/bash
# round-2 GPU call 9 (4 GPUs): strong-scaling line at N = 4 after the window rule (a window >= 4.5 tile layers), + the 2-rank tests
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 4 --steps 20 --warmup 3 --one-mode > gpurun_out/c9_bench512_n4.json 2> gpurun_out/c9_bench512_n4.err
echo "rc=$?" >> gpurun_out/c9_bench512_n4.err
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q 2>&1 | tail -5 > gpurun_out/c9_tests.log
tail -3 gpurun_out/c9_bench512_n4.err; cat gpurun_out/c9_tests.log
