#!/bin/bash
# One parametrised runner for everything that needs the B200 box (replaces the 47 one-off scripts of round 1).
#   gpurun --timeout 1500 -- 'bash scripts/gpu_run.sh <tag> <stage> [<stage> ...]'
# Outputs go to gpurun_out/<stage>_<tag>.*; copy what should be judged into profiles/ (named r02_*).
# Stages (at most ONE of the ncu_* / san_* families per call, B200_PROFILING.md):
#   tests            pytest -m gpu (the parity suite through the C ABI) + smoke()
#   bench            bench.py default line (N = 1) and the reference arm
#   bench32          bench.py --dtype f32
#   strong1          the headline kernel on ONE GPU at the per-GPU batch of a strong-scaling split over 2 / 4 / 8 GPUs
#   configs          scripts/bench_configs.py, fp64 and fp32
#   margin           scripts/parity_margin.py: fp64 / fp32 errors against the oracle at 20, 100 and 300 s
#   sweep            examples/sensitivity/truncation_sweep.py (15 150 points)
#   probe            scripts/probes/ffma_operands (built here: nvcc -O3 -o scripts/probes/ffma_operands scripts/probes/ffma_operands.cu)
#   variants         every cocofest_b200/csrc/build/var_*/libfes_b200.so (scripts/build_variants.sh) through
#                    scripts/bench_configs.py --configs $VCFG --dtype $VDT (defaults 2x0, f32), 60 s protocol
#   scale<N>         torchrun bench.py --gpus N (needs gpurun --gpus N)
#   ncu_list         launch list of the default bench command (after a plain run of it)
#   hess             tests/test_second_order_kernel.py on the GPU + config 5h (exact Hessian blocks)
#   ncu_full:<cfg>   ncu --set full of the top kernel of scripts/bench_configs.py --configs <cfg>  (2x0 = headline kernel)
#   san_memcheck | san_racecheck | san_synccheck   compute-sanitizer on the small parity problems (scripts/sanitize_cases.py)
set -u
tag=$1; shift
mkdir -p gpurun_out
for stage in "$@"; do
  echo "=== $stage ($tag) $(date +%T)"
  case $stage in
    tests)
      timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
      tail -5 gpurun_out/pytest_gpu_$tag.log
      timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_$tag.log
      tail -2 gpurun_out/smoke_$tag.log ;;
    bench)
      timeout 600 python bench.py --impl reference > gpurun_out/bench_reference_$tag.json 2> gpurun_out/bench_reference_$tag.err; echo "reference rc=$?"
      timeout 1200 python bench.py > gpurun_out/bench_n1_$tag.json 2> gpurun_out/bench_n1_$tag.err; echo "bench rc=$?"
      tail -3 gpurun_out/bench_n1_$tag.err; cut -c1-600 gpurun_out/bench_n1_$tag.json ;;
    bench32)
      timeout 900 python bench.py --dtype f32 --no-configs --no-cpu-baseline > gpurun_out/bench_n1_f32_$tag.json 2> gpurun_out/bench_n1_f32_$tag.err; echo "bench32 rc=$?"
      cut -c1-400 gpurun_out/bench_n1_f32_$tag.json ;;
    strong1)
      for b in 524288 262144 131072 65536 16384 4096 1024; do
        timeout 600 python bench.py --batch $b --steps 4 --no-configs --no-cpu-baseline --no-e2e 2>> gpurun_out/strong1_$tag.err |
          python -c "import sys, json; d = json.loads(sys.stdin.read()); print(json.dumps({'batch': d['config']['batch_per_gpu'], 'value': d['value'], 'ms': d['ms_per_step'], 'pipe': d['roofline']['pipe']['frac_of_nominal'] if d['roofline']['pipe'] else None}))" | tee -a gpurun_out/strong1_$tag.jsonl
      done ;;
    configs)
      timeout 900 python scripts/bench_configs.py --configs 1,2fwd,3,4,4b,4c,5,5c,5h,5pf > gpurun_out/configs_f64_$tag.jsonl 2> gpurun_out/configs_f64_$tag.err; echo "rc=$?"
      timeout 900 python scripts/bench_configs.py --configs 1,2fwd,2x0,3,4,4b,4c,5,5c,5pf --dtype f32 > gpurun_out/configs_f32_$tag.jsonl 2> gpurun_out/configs_f32_$tag.err; echo "rc=$?"
      cut -c1-260 gpurun_out/configs_f64_$tag.jsonl gpurun_out/configs_f32_$tag.jsonl ;;
    hess)
      timeout 600 python -m pytest tests/test_second_order_kernel.py tests/test_second_derivatives.py tests/test_api_surface.py -m gpu -q > gpurun_out/pytest_hess_$tag.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_hess_$tag.log
      tail -4 gpurun_out/pytest_hess_$tag.log
      timeout 600 python scripts/bench_configs.py --configs 5,5h > gpurun_out/configs_hess_$tag.jsonl 2> gpurun_out/configs_hess_$tag.err; echo "rc=$?"
      cut -c1-900 gpurun_out/configs_hess_$tag.jsonl ;;
    margin)
      timeout 900 python scripts/parity_margin.py > gpurun_out/parity_margin_$tag.txt 2>&1; echo "margin rc=$?"; cat gpurun_out/parity_margin_$tag.txt ;;
    sweep)
      timeout 1200 python examples/sensitivity/truncation_sweep.py --out-dir gpurun_out > gpurun_out/truncation_sweep_$tag.txt 2>&1; echo "sweep rc=$?"
      rm -f gpurun_out/truncation_*.pkl; cat gpurun_out/truncation_sweep_$tag.txt ;;
    probe)
      timeout 300 scripts/probes/ffma_operands > gpurun_out/probe_ffma_operands_$tag.txt 2>&1; echo "probe rc=$?"; cat gpurun_out/probe_ffma_operands_$tag.txt ;;
    variants)
      for lib in cocofest_b200/csrc/libfes_b200.so cocofest_b200/csrc/build/var_*/libfes_b200.so; do
        echo "## $lib" | tee -a gpurun_out/variants_$tag.txt
        FES_B200_LIB=$PWD/$lib timeout 600 python scripts/bench_configs.py --configs ${VCFG:-2x0} --dtype ${VDT:-f32} --t2 60 2>> gpurun_out/variants_$tag.err |
          python -c "import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if 'config' in d: print(d['config'][:50], d['dtype'], '%.4e' % d['value'], '%.3f ms' % d['ms'], 'regs', (d.get('roofline') or {}).get('registers'))" | tee -a gpurun_out/variants_$tag.txt
      done ;;
    scale*)
      n=${stage#scale}
      timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
        bench.py --gpus $n > gpurun_out/bench_n${n}_$tag.json 2> gpurun_out/bench_n${n}_$tag.err; echo "scale$n rc=$?"
      tail -2 gpurun_out/bench_n${n}_$tag.err; cut -c1-400 gpurun_out/bench_n${n}_$tag.json ;;
    ncu_list)
      cmd="python bench.py --steps 2 --warmup 3 --no-configs --no-cpu-baseline --no-e2e"
      timeout 600 $cmd > gpurun_out/ncu_list_plain_$tag.log 2>&1 &&
      timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_$tag.csv $cmd > gpurun_out/ncu_list_$tag.log 2>&1
      echo "ncu_list rc=$?"; tail -3 gpurun_out/launches_bench_$tag.csv ;;
    ncu_full:*)
      cfg=${stage#ncu_full:}; dt=f64; case $cfg in *@f32) dt=f32; cfg=${cfg%@f32};; esac
      pat=fes_ivp_kernel; case $cfg in 4b|4c) pat=fes_ident_kernel;; 5|5c|5pf) pat=fes_shoot_kernel;; 5h) pat=fes_hess_kernel;; esac
      cmd="python scripts/bench_configs.py --configs $cfg --dtype $dt --steps 1 --warmup 1 --t2 ${T2:-300}"
      timeout 600 $cmd > gpurun_out/ncu_full_plain_${cfg}_${dt}_$tag.log 2>&1 &&
      timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$pat -s 1 -c 1 -f -o gpurun_out/prof_${cfg}_${dt}_$tag $cmd > gpurun_out/ncu_full_${cfg}_${dt}_$tag.log 2>&1
      echo "ncu_full $cfg rc=$?"; tail -2 gpurun_out/ncu_full_${cfg}_${dt}_$tag.log ;;
    san_*)
      tool=${stage#san_}
      timeout 1500 compute-sanitizer --tool $tool --error-exitcode 3 python scripts/sanitize_cases.py > gpurun_out/sanitizer_${tool}_$tag.log 2>&1
      echo "sanitizer $tool rc=$?"; tail -6 gpurun_out/sanitizer_${tool}_$tag.log ;;
    *) echo "unknown stage $stage" ;;
  esac
done
echo "=== done $(date +%T)"
