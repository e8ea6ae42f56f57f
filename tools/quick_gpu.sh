#!/bin/bash
# GPU regression + kernel timing of the current build (run under gpurun)
timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 300 python bench.py --cells ${CELLS:-128} --steps 5 --warmup 3 --no-cpu --no-e2e "$@" 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value %.3e kernel-only %.3e' % (d['value'], d['roofline']['particle_steps_per_s_kernel_only']))
print({k: round(v,3) for k,v in d['phases_ms_per_step'].items()})"
