#!/bin/bash
# fourth GPU pass: identify parity, reciprocal-shortening margin, launch-shape sweep, configs incl. fes_identify
set -x
mkdir -p gpurun_out
V=cocofest_b200/csrc/build
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_d.log
tail -3 gpurun_out/pytest_gpu_d.log
{
  echo "== default build"; timeout 300 python scripts/parity_margin.py --seconds 20
  echo "== FES_RCP_SHORT build"; FES_B200_LIB=$PWD/$V/var_k_b64m5rcp/libfes_b200.so timeout 300 python scripts/parity_margin.py --seconds 20
  FES_B200_LIB=$PWD/$V/var_k_b64m5rcp/libfes_b200.so timeout 600 python -m pytest tests/test_parity_ivp.py tests/test_parity_shoot.py -m gpu -q 2>&1 | tail -3
} > gpurun_out/rcp_short_margin.txt 2>&1
cat gpurun_out/rcp_short_margin.txt
: > gpurun_out/tune_ivp_d.txt
for pass in 1 2; do
for d in $V/var_*/; do
  name=$(basename $d)
  out=$(FES_B200_LIB=$PWD/$d/libfes_b200.so timeout 300 python bench.py --steps 2 --warmup 3 --batch 1048576 --final-time 10 --no-cpu-baseline --no-e2e 2>/dev/null)
  echo "$name $(echo "$out" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["kernel_ms"], d["clocks"]["sm_mhz"])')" >> gpurun_out/tune_ivp_d.txt
done
done
sort gpurun_out/tune_ivp_d.txt
timeout 600 python scripts/bench_configs.py > gpurun_out/bench_configs_d.jsonl 2> gpurun_out/bench_configs_d.err; echo "rc=$?"
cat gpurun_out/bench_configs_d.jsonl; tail -5 gpurun_out/bench_configs_d.err
