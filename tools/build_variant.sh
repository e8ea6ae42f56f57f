#!/bin/bash
# build a library variant in a scratch dir: tools/build_variant.sh NAME "-DEB_T3X=8 ..."
set -e
name=$1; flags=$2
src=/root/repo/epoch_b200/csrc; out=/root/repo/gpurun_variants/$name
mkdir -p $out && cp $src/*.cu $src/*.cuh $src/Makefile $out/
sed -i 's#"../../include/epoch_b200.h"#"/root/repo/include/epoch_b200.h"#' $out/ctx.cuh
sed -i 's#../../include/epoch_b200.h#/root/repo/include/epoch_b200.h#' $out/Makefile
make -C $out -j8 EXTRA="$flags" 2>&1 | grep -E "error" || true
grep -A2 "k_push_splitILi3ELi1" $out/push_tiled.o.ptxas.log | grep -E "registers|spill" | tr '\n' ' '; echo " <- $name"
rm -f $out/*.o $out/*.cu $out/*.cuh
