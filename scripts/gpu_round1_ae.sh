#!/bin/bash
# ncu --set full of the identify kernel (config 4c), the Hmed ivp kernel (config 4) and the shoot kernel (config 5), final build
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_ident_kernel -s 9 -c 1 -f -o gpurun_out/prof_ident_cfg4_ae python scripts/bench_configs.py --configs 4 > gpurun_out/ncu_ident_ae.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_cfg4_ae python scripts/bench_configs.py --configs 4 > gpurun_out/ncu_ivp4_ae.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_shoot_kernel -s 3 -c 1 -f -o gpurun_out/prof_shoot_cfg5_ae python scripts/bench_configs.py --configs 5 > gpurun_out/ncu_shoot_ae.log 2>&1
timeout 600 python bench.py --dtype f32 --steps 2 --warmup 3 > gpurun_out/bench_f32_ae.json 2> gpurun_out/bench_f32_ae.err
ls gpurun_out | grep _ae
