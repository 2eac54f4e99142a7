#!/bin/bash
# tenth GPU pass: the whole GPU suite (no -x) incl. the bioptim-seam and edge-size tests
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/pytest_gpu_j.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_j.log
tail -12 gpurun_out/pytest_gpu_j.log
