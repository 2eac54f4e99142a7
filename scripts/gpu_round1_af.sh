#!/bin/bash
# propagator form of the two-state F-rows (FES_PROPAGATOR) against the default: parity on the GPU, configs 3 / 4
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
FES_B200_LIB=$PWD/$V/var_af_prop/libfes_b200.so timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_af.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_af.log
tail -4 gpurun_out/pytest_gpu_af.log
: > gpurun_out/tune_af.txt
for pass in 1 2; do
for d in default $V/var_af_*/; do
  if [ "$d" = default ]; then name=default; lib=$PWD/cocofest_b200/csrc/libfes_b200.so; else name=$(basename $d); lib=$PWD/$d/libfes_b200.so; fi
  FES_B200_LIB=$lib timeout 300 python scripts/bench_configs.py --configs 3,4 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name'", d["config"][:3], "%.4e" % d["value"], "%.4f" % d["ms_per_step"])' >> gpurun_out/tune_af.txt
done
done
sort gpurun_out/tune_af.txt
