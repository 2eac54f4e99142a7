#!/bin/bash
# ncu --set full of the two-state kernels of the final build (scalar step propagator): config 3 ivp, config 4 ivp, config 4 identify
set -x
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_cfg3_ak python scripts/bench_configs.py --configs 3 > gpurun_out/ncu_ivp3_ak.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_cfg4_ak python scripts/bench_configs.py --configs 4 > gpurun_out/ncu_ivp4_ak.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fes_ident_kernel -s 9 -c 1 -f -o gpurun_out/prof_ident_cfg4_ak python scripts/bench_configs.py --configs 4 > gpurun_out/ncu_ident_ak.log 2>&1
ls -la gpurun_out | grep _ak
