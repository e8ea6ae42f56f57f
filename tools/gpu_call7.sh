#!/bin/bash
# round-2 GPU call 7 (one B200): regression with the tiled 1D kernels and the thermal walls
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/c7_tests.log
tail -15 gpurun_out/c7_tests.log
