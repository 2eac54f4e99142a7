#!/bin/bash
# third GPU pass: parity of the stage-major kernels, launch-shape sweep, configs 1/3/4/5
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_c.log
tail -3 gpurun_out/pytest_gpu_c.log
: > gpurun_out/tune_ivp_c.txt
for d in cocofest_b200/csrc/build/var_*/; do
  name=$(basename $d)
  out=$(FES_B200_LIB=$PWD/$d/libfes_b200.so timeout 300 python bench.py --steps 2 --warmup 3 --batch 1048576 --final-time 10 --no-cpu-baseline --no-e2e 2>/dev/null)
  echo "$name $(echo "$out" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["kernel_ms"], d["roofline"]["frac"], d["clocks"]["sm_mhz"])')" >> gpurun_out/tune_ivp_c.txt
done
cat gpurun_out/tune_ivp_c.txt
timeout 600 python scripts/bench_configs.py > gpurun_out/bench_configs_c.jsonl 2> gpurun_out/bench_configs_c.err; echo "rc=$?"
cat gpurun_out/bench_configs_c.jsonl; tail -5 gpurun_out/bench_configs_c.err
