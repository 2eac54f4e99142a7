#!/bin/bash
# seventeenth GPU pass: compute-sanitizer memcheck + racecheck over the parity tests, examples test
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_examples.py -m gpu -q > gpurun_out/pytest_examples_q.log 2>&1; tail -3 gpurun_out/pytest_examples_q.log
CS=/usr/local/cuda/bin/compute-sanitizer
timeout 1200 $CS --tool memcheck --error-exitcode 7 --log-file gpurun_out/memcheck_q.log \
  python -m pytest tests/test_parity_ivp.py tests/test_parity_shoot.py tests/test_identify.py tests/test_bioptim_seam.py -m gpu -q -x \
  > gpurun_out/memcheck_pytest_q.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/memcheck_pytest_q.log
tail -3 gpurun_out/memcheck_pytest_q.log; tail -5 gpurun_out/memcheck_q.log
timeout 900 $CS --tool racecheck --error-exitcode 7 --log-file gpurun_out/racecheck_q.log \
  python -m pytest tests/test_parity_ivp.py tests/test_identify.py -m gpu -q -x -k "long_fatigue or forward_matches or cost_gradient or short_iteration" \
  > gpurun_out/racecheck_pytest_q.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/racecheck_pytest_q.log
tail -3 gpurun_out/racecheck_pytest_q.log; tail -5 gpurun_out/racecheck_q.log
