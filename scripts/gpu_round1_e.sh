#!/bin/bash
# fifth GPU pass: block-64 default build -- parity, smoke, full bench, configs (f64 + f32), ncu launch list and
# --set full captures of the three hot kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_e.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_e.log
tail -3 gpurun_out/pytest_gpu_e.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_e.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_e.log
tail -2 gpurun_out/smoke_e.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_full_e.json 2> gpurun_out/bench_full_e.err; echo "rc=$?"
cat gpurun_out/bench_full_e.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 3 > gpurun_out/bench_reference_e.json 2>&1
timeout 600 python scripts/bench_configs.py > gpurun_out/bench_configs_e.jsonl 2> gpurun_out/bench_configs_e.err
timeout 600 python scripts/bench_configs.py --dtype f32 > gpurun_out/bench_configs_e_f32.jsonl 2>> gpurun_out/bench_configs_e.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_e.csv $CMD > gpurun_out/ncu_list_e.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_x0_full $CMD > gpurun_out/ncu_full_e.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_shoot_kernel -s 3 -c 1 -f -o gpurun_out/prof_shoot_cfg5 python scripts/bench_configs.py --configs 5 > gpurun_out/ncu_shoot_e.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_ident_kernel -s 9 -c 1 -f -o gpurun_out/prof_ident_cfg4 python scripts/bench_configs.py --configs 4 > gpurun_out/ncu_ident_e.log 2>&1
ls -la gpurun_out
