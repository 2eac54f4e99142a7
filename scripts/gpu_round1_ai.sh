#!/bin/bash
# validation of HEAD (PROP2, 6 resident blocks): GPU suite, smoke, parity margin, both bench arms, f32 line, launch list,
# ncu --set full of the headline kernel and of the config-5 shoot kernel
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_ai.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_ai.log
tail -n 4 gpurun_out/pytest_gpu_ai.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_ai.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_ai.log; tail -n 2 gpurun_out/smoke_ai.log
timeout 300 python scripts/parity_margin.py --seconds 20 > gpurun_out/parity_margin_ai.txt 2>&1; cat gpurun_out/parity_margin_ai.txt
timeout 300 python bench.py --impl reference > gpurun_out/bench_reference_ai.json 2>/dev/null; echo "rc=$?"
timeout 900 python bench.py > gpurun_out/bench_default_ai.json 2> gpurun_out/bench_default_ai.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_default_ai.json
timeout 600 python bench.py --dtype f32 --steps 2 --warmup 3 > gpurun_out/bench_f32_ai.json 2> gpurun_out/bench_f32_ai.err; echo "rc=$?"
timeout 300 python scripts/bench_configs.py --configs 5 > gpurun_out/bench_configs_ai.jsonl 2> gpurun_out/bench_configs_ai.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_ai.csv $CMD > gpurun_out/ncu_list_ai.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_x0_full_ai $CMD > gpurun_out/ncu_full_ai.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fes_shoot_kernel -s 3 -c 1 -f -o gpurun_out/prof_shoot_cfg5_ai python scripts/bench_configs.py --configs 5 > gpurun_out/ncu_shoot_ai.log 2>&1
ls -la gpurun_out | grep _ai
