#!/bin/bash
# round-2 GPU call 15 (2 GPUs): smoke run of tools/run_configs_mgpu.py at 1/8 scale
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29761 tools/run_configs_mgpu.py both 0.125 > gpurun_out/c15_cfg_n2.jsonl 2> gpurun_out/c15_cfg_n2.err
echo "rc=$?" >> gpurun_out/c15_cfg_n2.err
tail -15 gpurun_out/c15_cfg_n2.err | cut -c1-300; cat gpurun_out/c15_cfg_n2.jsonl | cut -c1-1500
