#!/bin/bash
# round-2 GPU call 1: regression, kernel variants at 512^3, ncu full capture at 128^3
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/c1_tests.log
for v in main base pair bal; do
  lib=$PWD/gpurun_variants/$v/libepoch_b200.so
  [ $v = main ] && lib=$PWD/epoch_b200/csrc/libepoch_b200.so
  EB200_LIB=$lib timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e 2> gpurun_out/c1_bench_$v.err | tail -1 > gpurun_out/c1_bench_$v.json
done
timeout 500 ncu --set full --import-source on --clock-control none -k regex:'k_push_split|k_deposit_split' -s 8 -c 4 -f -o gpurun_out/r02_pairbal_128 \
  python bench.py --cells 128 --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/c1_ncu.log 2>&1
ls -la gpurun_out
