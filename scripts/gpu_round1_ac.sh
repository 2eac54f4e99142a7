#!/bin/bash
# streaming (evict-first) trace stores against plain stores: configs 1 and 3 (trace-heavy), two passes
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
: > gpurun_out/tune_ac.txt
for pass in 1 2; do
for d in default $V/var_ac_*/; do
  if [ "$d" = default ]; then name=default; lib=$PWD/cocofest_b200/csrc/libfes_b200.so; else name=$(basename $d); lib=$PWD/$d/libfes_b200.so; fi
  FES_B200_LIB=$lib timeout 300 python scripts/bench_configs.py --configs 1,3,4 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name'", d["config"][:3], "%.4e" % d["value"], "%.4f" % d["ms_per_step"])' >> gpurun_out/tune_ac.txt
done
done
sort gpurun_out/tune_ac.txt
