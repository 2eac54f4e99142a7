#!/bin/bash
# final validation of the round: GPU suite, smoke, both bench arms
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_final.log
tail -4 gpurun_out/pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_final.log; tail -2 gpurun_out/smoke_final.log
timeout 300 python bench.py --impl reference > gpurun_out/bench_reference_final.json 2>/dev/null; echo "rc=$?"
timeout 900 python bench.py > gpurun_out/bench_default_final.json 2> gpurun_out/bench_default_final.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_default_final.json
