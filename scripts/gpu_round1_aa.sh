#!/bin/bash
# ncu of the shoot kernel (config 5) after the stage-table change
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_shoot_kernel -s 3 -c 1 -f -o gpurun_out/prof_shoot_cfg5_aa python scripts/bench_configs.py --configs 5 > gpurun_out/ncu_shoot_aa.log 2>&1
tail -2 gpurun_out/ncu_shoot_aa.log
