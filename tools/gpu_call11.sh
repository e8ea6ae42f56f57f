#!/bin/bash
# round-2 GPU call 11 (one B200): final regression, the bench line, the other BASELINE configs at full size
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/c11_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/c11_bench.json 2> gpurun_out/c11_bench.err
echo "rc=$?" >> gpurun_out/c11_bench.err
timeout 240 python tools/run_configs.py cfg2 > gpurun_out/c11_configs.jsonl 2> gpurun_out/c11_configs.err
timeout 300 python tools/run_configs.py cfg4 >> gpurun_out/c11_configs.jsonl 2>> gpurun_out/c11_configs.err
timeout 120 python tools/run_configs.py cfg5 >> gpurun_out/c11_configs.jsonl 2>> gpurun_out/c11_configs.err
cat gpurun_out/c11_tests.log; tail -2 gpurun_out/c11_bench.err; wc -l gpurun_out/c11_configs.jsonl
