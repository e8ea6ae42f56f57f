#!/bin/bash
# run bench (kernel-only figure) for every library variant under gpurun_variants/
for d in gpurun_variants/*/; do
  n=$(basename $d)
  EB200_LIB=$PWD/$d/libepoch_b200.so timeout 300 python bench.py --cells ${CELLS:-128} --steps 5 --warmup 3 --no-cpu --no-e2e 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$n', '%.3e'%d['value'], '%.3e'%d['roofline']['particle_steps_per_s_kernel_only'], 'push_ms %.3f'%d['phases_ms_per_step']['push_kernel'], 'A %.3f B %.3f'%(d['phases_ms_per_step']['push_A'], d['phases_ms_per_step']['push_B']))"
done
