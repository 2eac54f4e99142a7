#!/bin/bash
# twenty-fifth GPU pass: stage table with the psi modes (nominal Cn and forced Cn columns without stages of their own)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_y.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_y.log
tail -4 gpurun_out/pytest_gpu_y.log
timeout 300 python scripts/parity_margin.py --seconds 20 > gpurun_out/parity_margin_y.txt 2>&1; cat gpurun_out/parity_margin_y.txt
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 > gpurun_out/bench_configs_y.jsonl 2> gpurun_out/bench_configs_y.err
python -c '
import json
for l in open("gpurun_out/bench_configs_y.jsonl"):
    d = json.loads(l)
    if "config" in d: print(d["config"][:3], "%.4e" % d["value"], round(d["ms_per_step"], 4))'
timeout 900 python bench.py > gpurun_out/bench_default_y.json 2> gpurun_out/bench_default_y.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_default_y.json
