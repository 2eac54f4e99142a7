#!/bin/bash
# Builds launch-shape variants of libfes_b200.so for the on-GPU sweep (scripts/tune_ivp.sh).
# name:block:min_blocks(0 = unspecified):maxrregcount(0 = none)[:extra define]
set -e
cd "$(dirname "$0")/../cocofest_b200/csrc"
for v in ${VARIANTS:-b128m0:128:0:0 b128m1:128:1:0 b128m3:128:3:0 b64m5:64:5:0 b64m6:64:6:0 b256m1:256:1:0 b96m4:96:4:0 b128m4:128:4:0}; do
  IFS=: read name blk minb maxr defs <<< "$v"
  extra="-DFES_BLOCK=$blk"
  [ "$minb" != "0" ] && extra="$extra -DFES_MIN_BLOCKS=$minb"
  [ "$maxr" != "0" ] && extra="$extra -maxrregcount=$maxr"
  [ -n "$defs" ] && extra="$extra -D$defs"
  make -j8 BUILD=build/var_$name LIB=build/var_$name/libfes_b200.so COUNTS=build/var_$name/sass_counts.json EXTRA="$extra" > build/var_$name.log 2>&1 || { tail -5 build/var_$name.log; exit 1; }
  echo "$name: $(grep -A3 'fes_ivp_kernelIdLi0ELb1ELi1E' build/var_$name/fes_inst_k0f1.ptxas.log | grep -o 'Used [0-9]* registers\|[0-9]* bytes spill stores' | tr '\n' ' ')"
done
