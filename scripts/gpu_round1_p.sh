#!/bin/bash
# sixteenth GPU pass: stream / CUDA-graph / NaN tests
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_streams_and_nan.py tests/test_c_abi.py -m gpu -q > gpurun_out/pytest_gpu_p.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_p.log
tail -30 gpurun_out/pytest_gpu_p.log | cut -c1-220
