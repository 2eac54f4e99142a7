#!/bin/bash
# fifteenth GPU pass: final validation of HEAD -- GPU suite, smoke, default bench (both arms)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_o.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_o.log
tail -4 gpurun_out/pytest_gpu_o.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_o.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_o.log
tail -2 gpurun_out/smoke_o.log
timeout 300 python bench.py --impl reference > gpurun_out/bench_reference_o.json 2> gpurun_out/bench_reference_o.err; echo "rc=$?"
timeout 900 python bench.py > gpurun_out/bench_default_o.json 2> gpurun_out/bench_default_o.err; echo "rc=$?"
cut -c1-300 gpurun_out/bench_default_o.json
