#!/bin/bash
# 2 x 2 step propagator of the reduced fatigue columns (FES_PROP2) against the default: parity on the GPU, headline, configs
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
P2=$PWD/$V/var_ag_prop2/libfes_b200.so
FES_B200_LIB=$P2 timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_ag.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_ag.log
tail -n 4 gpurun_out/pytest_gpu_ag.log
FES_B200_LIB=$P2 timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_n1_ag_prop2.json 2> gpurun_out/bench_n1_ag_prop2.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_n1_ag_prop2.json
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_n1_ag_default.json 2> gpurun_out/bench_n1_ag_default.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_n1_ag_default.json
: > gpurun_out/tune_ag.txt
for pass in 1 2; do
for name in default var_ag_prop2; do
  if [ "$name" = default ]; then lib=$PWD/cocofest_b200/csrc/libfes_b200.so; else lib=$PWD/$V/$name/libfes_b200.so; fi
  for dt in f64 f32; do
  FES_B200_LIB=$lib timeout 300 python scripts/bench_configs.py --configs 5 --dtype $dt 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name' '$dt'", d["config"][:3], "%.4e" % d["value"], "%.4f" % d["ms_per_step"])' >> gpurun_out/tune_ag.txt
  done
done
done
FES_B200_LIB=$P2 timeout 600 python bench.py --dtype f32 --steps 2 --warmup 3 > gpurun_out/bench_n1_f32_ag_prop2.json 2> gpurun_out/bench_n1_f32_ag_prop2.err; echo "rc=$?"
sort gpurun_out/tune_ag.txt
