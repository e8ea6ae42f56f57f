#!/bin/bash
# round-2 GPU call 10 (2 GPUs): NCCL p2p settings against the halo phase of the strong-scaling run (512^3 cut in two)
mkdir -p gpurun_out
run() {   # name, env...
  name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 \
    bench.py --gpus 2 --steps 10 --warmup 3 --one-mode $E2E > gpurun_out/c10_$name.json 2> gpurun_out/c10_$name.err
  echo "rc=$?" >> gpurun_out/c10_$name.err
}
E2E=""
run default NCCL_DEBUG=WARN
E2E="--no-e2e"
run minch16 NCCL_MIN_P2P_NCHANNELS=16 NCCL_MAX_P2P_NCHANNELS=32
run cumemcpy NCCL_P2P_USE_CUDA_MEMCPY=1
for n in default minch16 cumemcpy; do tail -c 300 gpurun_out/c10_$n.err | tail -2; python - <<P
import json
try:
    d=json.loads(open('gpurun_out/c10_$n.json').read().strip().splitlines()[-1])
    print('$n', '%.3e'%d['value'], '%.2f ms'%d['ms_per_step'], {k:round(v,2) for k,v in d['phases_ms_per_step_max_over_ranks'].items()}, d['sorts_whole_run'])
except Exception as e: print('$n', 'failed', e)
P
done
