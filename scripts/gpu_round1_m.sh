#!/bin/bash
# thirteenth GPU pass: resident-blocks sweep (FES_MIN_BLOCKS) over configs 1, 2 (forward only, 30 s), 3, 4, 5; f32 headline
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
: > gpurun_out/tune_m.txt
for d in default $V/var_m_*/; do
  if [ "$d" = default ]; then name=default; lib=$PWD/cocofest_b200/csrc/libfes_b200.so; else name=$(basename $d); lib=$PWD/$d/libfes_b200.so; fi
  [ -f "$lib" ] || continue
  FES_B200_LIB=$lib timeout 300 python scripts/bench_configs.py --configs 1,2,3,4,5 --t2 30 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name'", d["config"][:2], "%.4e" % d["value"], "%.4f" % d["ms_per_step"])' >> gpurun_out/tune_m.txt
done
cat gpurun_out/tune_m.txt
timeout 600 python bench.py --dtype f32 --steps 2 --warmup 3 > gpurun_out/bench_f32_m.json 2> gpurun_out/bench_f32_m.err; echo "rc=$?"
cut -c1-400 gpurun_out/bench_f32_m.json
