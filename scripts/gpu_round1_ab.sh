#!/bin/bash
# validation of HEAD (mu / nu from stage factors): GPU suite, parity margin, configs, default bench, ncu of the headline kernel
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_ab.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_ab.log
tail -4 gpurun_out/pytest_gpu_ab.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_ab.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_ab.log; tail -2 gpurun_out/smoke_ab.log
timeout 300 python scripts/parity_margin.py --seconds 20 > gpurun_out/parity_margin_ab.txt 2>&1; cat gpurun_out/parity_margin_ab.txt
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 > gpurun_out/bench_configs_ab.jsonl 2> gpurun_out/bench_configs_ab.err
timeout 600 python scripts/bench_configs.py --configs 1,2,3,4,5 --dtype f32 > gpurun_out/bench_configs_ab_f32.jsonl 2>> gpurun_out/bench_configs_ab.err
timeout 900 python bench.py > gpurun_out/bench_default_ab.json 2> gpurun_out/bench_default_ab.err; echo "rc=$?"
cut -c1-200 gpurun_out/bench_default_ab.json
timeout 300 python bench.py --impl reference > gpurun_out/bench_reference_ab.json 2>/dev/null
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_ab.csv $CMD > gpurun_out/ncu_list_ab.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fes_ivp_kernel -s 3 -c 1 -f -o gpurun_out/prof_ivp_x0_full_ab $CMD > gpurun_out/ncu_full_ab.log 2>&1
ls gpurun_out | head -50
