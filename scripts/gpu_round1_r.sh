#!/bin/bash
# eighteenth GPU pass: HEAD after the Ding2007 P launch-bound change -- GPU suite, configs, default bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_r.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r.log
tail -4 gpurun_out/pytest_gpu_r.log
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 > gpurun_out/bench_configs_r.jsonl 2> gpurun_out/bench_configs_r.err
python -c '
import json
for l in open("gpurun_out/bench_configs_r.jsonl"):
    d = json.loads(l)
    if "config" in d: print(d["config"][:40], "%.4e" % d["value"], d["ms_per_step"])'
timeout 900 python bench.py > gpurun_out/bench_default_r.json 2> gpurun_out/bench_default_r.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_default_r.json
