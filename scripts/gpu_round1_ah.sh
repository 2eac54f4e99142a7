#!/bin/bash
# launch-shape sweep of the PROP2 build: headline kernel (10 s protocol) and config 5 per variant, then all configs of the default
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
: > gpurun_out/tune_ah.txt
for pass in 1 2; do
for name in default var_ah_m4 var_ah_m6 var_ah_b32m10 var_ah_b96m3 var_ah_b128m3 var_ah_m0; do
  if [ "$name" = default ]; then lib=$PWD/cocofest_b200/csrc/libfes_b200.so; else lib=$PWD/$V/$name/libfes_b200.so; fi
  out=$(FES_B200_LIB=$lib timeout 300 python bench.py --steps 2 --warmup 3 --batch 1048576 --final-time 10 --no-cpu-baseline --no-e2e 2>/dev/null)
  echo "$name headline $(echo "$out" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("%.4e" % d["value"], d["roofline"]["kernel_ms"], d["clocks"]["sm_mhz"])')" >> gpurun_out/tune_ah.txt
  FES_B200_LIB=$lib timeout 300 python scripts/bench_configs.py --configs 5 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name' cfg", d["config"][:3], "%.4e" % d["value"], "%.4f" % d["ms_per_step"])' >> gpurun_out/tune_ah.txt
done
done
sort gpurun_out/tune_ah.txt
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 > gpurun_out/bench_configs_ah.jsonl 2> gpurun_out/bench_configs_ah.err; echo "rc=$?"
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 --dtype f32 > gpurun_out/bench_configs_ah_f32.jsonl 2> gpurun_out/bench_configs_ah_f32.err; echo "rc=$?"
cut -c1-160 gpurun_out/bench_configs_ah.jsonl
