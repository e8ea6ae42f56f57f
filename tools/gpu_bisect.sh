#!/bin/bash
run() { echo "=== $1"; shift; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT tests/mgpu_worker.py 2 $STEPS triangle balance 2>&1 | grep -E "balance:|parity|pos 0 0|field ex|field jz|Error|error|re-slicing" | head -8; PORT=$((PORT+1)); }
PORT=29800
STEPS=6 run "plain" A=1
STEPS=3 run "3 steps (one step after the re-slice)" A=1
