#!/bin/bash
# fourteenth GPU pass: ncu --set full of the f32 headline kernel (short protocol)
set -x
mkdir -p gpurun_out
CMD="python bench.py --dtype f32 --steps 2 --warmup 3 --final-time 30 --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_x0_f32_n $CMD > gpurun_out/ncu_f32_n.log 2>&1
tail -3 gpurun_out/ncu_f32_n.log
