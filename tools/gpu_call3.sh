#!/bin/bash
# round-2 GPU call 3 (2 GPUs): the no-alloc test alone in a fresh process, then the whole GPU suite (no -x: see every failure)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15 > gpurun_out/c3_zz_alone.log
timeout 1800 python -m pytest tests -m gpu -q --durations=10 2>&1 | tail -60 > gpurun_out/c3_tests.log
tail -3 gpurun_out/c3_zz_alone.log gpurun_out/c3_tests.log
