#!/bin/bash
# weak scaling of the headline bench with the PROP2 build at N = 2, 4 on one box (gpurun --gpus 4; N = 1 is pass ai)
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/scale_aj_gpus.txt 2>&1
P=29610
for N in 2 4; do
  P=$((P+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P \
    bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/bench_n${N}_aj.json 2> gpurun_out/bench_n${N}_aj.err; echo "n$N rc=$?"
done
for N in 2 4; do python - $N <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/bench_n{n}_aj.json") if l.startswith("{")][-1])
    print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"])
except Exception as e:
    print(n, "failed", e)
PY
done
