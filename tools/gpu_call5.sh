#!/bin/bash
# round-2 GPU call 5 (one B200): regression after the k_src / field-grid changes, variants, push_overlap option, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/c5_tests.log
timeout 700 python tools/variant_sweep.py --cells 512 --steps 5 --warmup 3 --check main=default sub1=gpurun_variants/sub1/libepoch_b200.so \
  srcu8=gpurun_variants/srcu8/libepoch_b200.so > gpurun_out/c5_variants.jsonl 2> gpurun_out/c5_variants.err
timeout 300 python tools/variant_sweep.py --cells 512 --steps 5 --warmup 3 --check --opt push_overlap=1 overlap=default >> gpurun_out/c5_variants.jsonl 2>> gpurun_out/c5_variants.err
timeout 400 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:'k_update_e|k_update_b|k_src' -s 10 -c 10 --csv --log-file gpurun_out/r02_dram_512_fields_after.csv \
  python tools/variant_sweep.py --name ncu --cells 512 --steps 2 --warmup 2 > gpurun_out/c5_ncu_dram.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^(?!k_unpack_aos)' -c 400 --csv --log-file gpurun_out/r02_launches_512.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-check > gpurun_out/c5_ncu_launch.log 2>&1
cat gpurun_out/c5_tests.log gpurun_out/c5_variants.jsonl
