#!/bin/bash
# twelfth GPU pass: ncu --set full captures of the ivp kernel in configs 1 (HBM-write heavy), 3 and 4 (parameter columns)
set -x
mkdir -p gpurun_out
for c in 1 3 4; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f \
    -o gpurun_out/prof_ivp_cfg${c}_l python scripts/bench_configs.py --configs $c > gpurun_out/ncu_cfg${c}_l.log 2>&1
done
ls -la gpurun_out
