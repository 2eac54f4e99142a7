#!/bin/bash
# ninth GPU pass: HEAD (reduced nominal fatigue rows, identify scratch in shared memory) -- parity incl. full-size
# tests, smoke, full bench + reference arm, configs (f64 + f32), ncu launch list and --set full captures
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu_i.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_i.log
tail -15 gpurun_out/pytest_gpu_i.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_i.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_i.log
tail -2 gpurun_out/smoke_i.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_full_i.json 2> gpurun_out/bench_full_i.err; echo "rc=$?"
cat gpurun_out/bench_full_i.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 3 > gpurun_out/bench_reference_i.json 2>&1
timeout 300 python scripts/parity_margin.py --seconds 20 > gpurun_out/parity_margin_i.txt 2>&1; cat gpurun_out/parity_margin_i.txt
timeout 600 python scripts/bench_configs.py > gpurun_out/bench_configs_i.jsonl 2> gpurun_out/bench_configs_i.err
timeout 600 python scripts/bench_configs.py --dtype f32 > gpurun_out/bench_configs_i_f32.jsonl 2>> gpurun_out/bench_configs_i.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_i.csv $CMD > gpurun_out/ncu_list_i.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_x0_full_i $CMD > gpurun_out/ncu_full_i.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_shoot_kernel -s 3 -c 1 -f -o gpurun_out/prof_shoot_cfg5_i python scripts/bench_configs.py --configs 5 > gpurun_out/ncu_shoot_i.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_ident_kernel -s 9 -c 1 -f -o gpurun_out/prof_ident_cfg4_i python scripts/bench_configs.py --configs 4 > gpurun_out/ncu_ident_i.log 2>&1
ls -la gpurun_out
