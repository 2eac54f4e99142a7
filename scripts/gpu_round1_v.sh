#!/bin/bash
# twenty-second GPU pass: shared-mode treatment of the forced Cn components (HEAD) against the previous build
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_v.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_v.log
tail -4 gpurun_out/pytest_gpu_v.log
V=cocofest_b200/csrc/build
: > gpurun_out/tune_v.txt
for d in default $V/var_v_*/; do
  if [ "$d" = default ]; then name=default; lib=$PWD/cocofest_b200/csrc/libfes_b200.so; else name=$(basename $d); lib=$PWD/$d/libfes_b200.so; fi
  FES_B200_LIB=$lib timeout 300 python scripts/bench_configs.py --configs 3,4,5 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name'", d["config"][:3], "%.4e" % d["value"], "%.4f" % d["ms_per_step"])' >> gpurun_out/tune_v.txt
done
cat gpurun_out/tune_v.txt
