#!/bin/bash
# round-2 GPU call 12 (one B200): the count-ranked deposit map on a THERMALISED plasma (20 warm-up steps: cell counts Poisson, not the load's uniform 16)
mkdir -p gpurun_out
timeout 500 python tools/variant_sweep.py --cells 512 --steps 5 --warmup 20 main=default bal=gpurun_variants/bal/libepoch_b200.so > gpurun_out/c12_variants.jsonl 2> gpurun_out/c12_variants.err
cat gpurun_out/c12_variants.jsonl
