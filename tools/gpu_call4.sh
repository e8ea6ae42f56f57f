#!/bin/bash
# round-2 GPU call 4 (one B200): GPU regression, push/deposit variants at 512^3, ncu captures on the bench config, bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/c4_tests.log
timeout 900 python tools/variant_sweep.py --cells 512 --steps 5 --warmup 3 main=default base=gpurun_variants/base/libepoch_b200.so \
  sub4=gpurun_variants/sub4/libepoch_b200.so minb3=gpurun_variants/minb3/libepoch_b200.so mina3=gpurun_variants/mina3/libepoch_b200.so \
  > gpurun_out/c4_variants.jsonl 2> gpurun_out/c4_variants.err
# ncu --set full: one z slab of the bench box (512 x 512 x 128 cells x 16 ppc: the state of one rank at N = 4; small enough
# for ncu's save / restore of the kernel's memory between its ~40 replay passes), one whole step (4 field + push / deposit launches)
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'k_push_split|k_deposit_split|k_update_e|k_update_b' -s 12 -c 10 -f \
  -o gpurun_out/r02_slab512 python tools/variant_sweep.py --name ncu --cells 512 --cells-z 128 --steps 2 --warmup 2 > gpurun_out/c4_ncu_slab.log 2>&1
# the full 512^3 state: DRAM bytes + duration only (single pass, no replay), one whole step
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:'k_push_split|k_deposit_split|k_update_e|k_update_b|k_src' -s 60 -c 60 --csv --log-file gpurun_out/r02_dram_512.csv \
  python tools/variant_sweep.py --name ncu --cells 512 --steps 2 --warmup 2 > gpurun_out/c4_ncu_dram.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/c4_bench.json 2> gpurun_out/c4_bench.err
echo "rc=$?" >> gpurun_out/c4_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_512.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-check > gpurun_out/c4_ncu_launch.log 2>&1
ls -la gpurun_out | tail -20
cat gpurun_out/c4_tests.log gpurun_out/c4_variants.jsonl
