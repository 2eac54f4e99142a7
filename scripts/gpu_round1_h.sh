#!/bin/bash
# eighth GPU pass: full-size tests, identify register-cap sweep (scratch in shared memory), configs, full bench
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu_h.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_h.log
tail -15 gpurun_out/pytest_gpu_h.log
: > gpurun_out/tune_ident_h.txt
for d in $V/var_h_id*/; do
  name=$(basename $d)
  FES_B200_LIB=$PWD/$d/libfes_b200.so timeout 300 python scripts/bench_configs.py --configs 4 2>/dev/null | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if "config" in d: print("'$name'", d["config"][:3], d["value"], d["ms_per_step"])' >> gpurun_out/tune_ident_h.txt
done
cat gpurun_out/tune_ident_h.txt
timeout 600 python scripts/bench_configs.py > gpurun_out/bench_configs_h.jsonl 2> gpurun_out/bench_configs_h.err
python -c '
import json
for l in open("gpurun_out/bench_configs_h.jsonl"):
    d = json.loads(l)
    if "config" in d: print(d["config"][:40], d["value"], d["ms_per_step"])'
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_full_h.json 2> gpurun_out/bench_full_h.err; echo "rc=$?"
cut -c1-300 gpurun_out/bench_full_h.json
