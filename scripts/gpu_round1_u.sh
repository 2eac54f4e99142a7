#!/bin/bash
# twenty-first GPU pass: HEAD with 6-block identify kernel and slow-path-free setup reciprocals -- GPU suite, configs, bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_u.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_u.log
tail -4 gpurun_out/pytest_gpu_u.log
timeout 300 python scripts/parity_margin.py --seconds 20 > gpurun_out/parity_margin_u.txt 2>&1; cat gpurun_out/parity_margin_u.txt
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 > gpurun_out/bench_configs_u.jsonl 2> gpurun_out/bench_configs_u.err
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 --dtype f32 > gpurun_out/bench_configs_u_f32.jsonl 2>> gpurun_out/bench_configs_u.err
python -c '
import json
for f in ("gpurun_out/bench_configs_u.jsonl", "gpurun_out/bench_configs_u_f32.jsonl"):
  for l in open(f):
    d = json.loads(l)
    if "config" in d: print(d["dtype"], d["config"][:40], "%.4e" % d["value"], round(d["ms_per_step"], 4))'
timeout 900 python bench.py > gpurun_out/bench_default_u.json 2> gpurun_out/bench_default_u.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_default_u.json
