for lib in cocofest_b200/csrc/libfes_b200.so cocofest_b200/csrc/build/var_*/libfes_b200.so; do
  echo "## $lib" | tee -a gpurun_out/latency_ac.txt
  for b in 1048576 65536 16384 1024; do
    FES_B200_LIB=$PWD/$lib timeout 600 python bench.py --batch $b --steps 3 --no-configs --no-cpu-baseline --no-e2e 2>> gpurun_out/latency_ac.err |
      python -c "import sys, json; d = json.loads(sys.stdin.read()); print(d['config']['batch_per_gpu'], '%.4e' % d['value'], '%.2f ms' % d['ms_per_step'])" | tee -a gpurun_out/latency_ac.txt
  done
done
