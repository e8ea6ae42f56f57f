#!/bin/bash
# round-2 GPU call 16 (8 GPUs): configs[4] (2D laser-foil, cut along y, re-balanced) and configs[3] (3D LWFA, bspline3, moving window, 1 x 2 x 4 pencils) at full size
mkdir -p gpurun_out
timeout 88 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29771 tools/run_configs_mgpu.py both 1.0 > gpurun_out/r02_configs_8gpu.jsonl 2> gpurun_out/c16_cfg_n8.err
echo "rc=$?" >> gpurun_out/c16_cfg_n8.err
tail -4 gpurun_out/c16_cfg_n8.err | cut -c1-300; cut -c1-900 gpurun_out/r02_configs_8gpu.jsonl
