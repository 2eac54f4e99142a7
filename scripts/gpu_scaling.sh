#!/bin/bash
# eleventh GPU pass (gpurun --gpus 8): weak scaling of the headline bench at N = 1, 2, 4, 8 on one box
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/scale_ad_gpus.txt 2>&1
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 > gpurun_out/bench_n1_ad.json 2> gpurun_out/bench_n1_ad.err; echo "n1 rc=$?"
P=29510
for N in 2 4 8; do
  P=$((P+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P \
    bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/bench_n${N}_ad.json 2> gpurun_out/bench_n${N}_ad.err; echo "n$N rc=$?"
done
for N in 1 2 4 8; do python - $N <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/bench_n{n}_ad.json") if l.startswith("{")][-1])
    print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"])
except Exception as e:
    print(n, "failed", e)
PY
done
