#!/bin/bash
# second GPU pass: parity of the reduced-tangent kernels, launch-shape sweep, full bench
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_b.log
tail -3 gpurun_out/pytest_gpu_b.log
: > gpurun_out/tune_ivp.txt
for d in cocofest_b200/csrc/build/var_*/; do
  name=$(basename $d)
  out=$(FES_B200_LIB=$PWD/$d/libfes_b200.so timeout 300 python bench.py --steps 2 --warmup 3 --batch 1048576 --final-time 10 --no-cpu-baseline --no-e2e 2>/dev/null)
  echo "$name $(echo "$out" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["kernel_ms"], d["roofline"]["frac"], d["clocks"]["sm_mhz"])')" >> gpurun_out/tune_ivp.txt
done
cat gpurun_out/tune_ivp.txt
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_full_b.json 2> gpurun_out/bench_full_b.err; echo "rc=$?"
cat gpurun_out/bench_full_b.json
