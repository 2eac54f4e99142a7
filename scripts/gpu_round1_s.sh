#!/bin/bash
# nineteenth GPU pass: nominal-ahead stage order (FES_NOMINAL_AHEAD) against the default, headline kernel, 30 s protocol
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
: > gpurun_out/tune_s.txt
for pass in 1 2; do
for d in default $V/var_s_*/; do
  if [ "$d" = default ]; then name=default; lib=$PWD/cocofest_b200/csrc/libfes_b200.so; else name=$(basename $d); lib=$PWD/$d/libfes_b200.so; fi
  out=$(FES_B200_LIB=$lib timeout 300 python bench.py --steps 3 --warmup 3 --final-time 30 --no-cpu-baseline --no-e2e 2>/dev/null)
  echo "$name $(echo "$out" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["kernel_ms"], d["clocks"]["sm_mhz"])')" >> gpurun_out/tune_s.txt
done
done
sort gpurun_out/tune_s.txt
