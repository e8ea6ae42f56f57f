#!/bin/bash
# usage: build_and_run.sh EPOCH_CHECKOUT OUTDIR [SHAPE_DEFINE]
# builds epoch1d/2d/3d with gfortran and produces the reference dumps consumed by tests/test_ref_dump.py
set -euo pipefail
EPOCH=$(realpath "$1"); OUT=$(realpath -m "$2"); DEF=${3:-}
HERE=$(cd "$(dirname "$0")" && pwd)
mkdir -p "$OUT"
( cd "$EPOCH" && git submodule update --init SDF )
for nd in 1 2 3; do
  ( cd "$EPOCH/epoch${nd}d" && make clean >/dev/null 2>&1 || true; make COMPILER=gfortran ${DEF:+DEFINES="$DEF"} -j8 )
  run="$OUT/run${nd}d"; rm -rf "$run"; mkdir -p "$run"
  cp "$HERE/decks/${nd}d/input.deck" "$run/"
  ( cd "$EPOCH/epoch${nd}d" && echo "$run" | mpirun -np 1 ./bin/epoch${nd}d )
  PYTHONPATH="$EPOCH/SDF/utilities:${PYTHONPATH:-}" python3 "$HERE/sdf2npz.py" "$run" "$OUT" "ref${nd}d"
done
ls -la "$OUT"/*.npz
