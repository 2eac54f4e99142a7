#!/bin/bash
# seventh GPU pass: nominal-fatigue-row reduction -- parity, launch-shape sweep (ivp) and register-cap sweep (identify)
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_g.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_g.log
tail -3 gpurun_out/pytest_gpu_g.log
timeout 300 python scripts/parity_margin.py --seconds 20 > gpurun_out/parity_margin_g.txt 2>&1; cat gpurun_out/parity_margin_g.txt
: > gpurun_out/tune_ivp_g.txt
for pass in 1 2; do
for d in $V/var_g_b*/; do
  name=$(basename $d)
  out=$(FES_B200_LIB=$PWD/$d/libfes_b200.so timeout 300 python bench.py --steps 2 --warmup 3 --batch 1048576 --final-time 10 --no-cpu-baseline --no-e2e 2>/dev/null)
  echo "$name $(echo "$out" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["kernel_ms"], d["clocks"]["sm_mhz"])')" >> gpurun_out/tune_ivp_g.txt
done
done
sort gpurun_out/tune_ivp_g.txt
: > gpurun_out/tune_ident_g.txt
for d in $V/var_g_b64m5/ $V/var_g_id*/; do
  name=$(basename $d)
  FES_B200_LIB=$PWD/$d/libfes_b200.so timeout 300 python scripts/bench_configs.py --configs 4 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name'", d["config"][:3], d["value"], d["ms_per_step"])' >> gpurun_out/tune_ident_g.txt
done
cat gpurun_out/tune_ident_g.txt
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_full_g.json 2> gpurun_out/bench_full_g.err; echo "rc=$?"
cut -c1-300 gpurun_out/bench_full_g.json
