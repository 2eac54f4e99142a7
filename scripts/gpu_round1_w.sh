#!/bin/bash
# twenty-third GPU pass: tanh cache of the Hmed window slots + gated shared Cn modes -- GPU suite, configs
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_w.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_w.log
tail -4 gpurun_out/pytest_gpu_w.log
timeout 600 python scripts/bench_configs.py --configs 3,4,5 > gpurun_out/bench_configs_w.jsonl 2> gpurun_out/bench_configs_w.err
python -c '
import json
for l in open("gpurun_out/bench_configs_w.jsonl"):
    d = json.loads(l)
    if "config" in d: print(d["config"][:3], "%.4e" % d["value"], round(d["ms_per_step"], 4))'
