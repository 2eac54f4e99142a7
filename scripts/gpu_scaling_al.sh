#!/bin/bash
# N = 8 line of the headline bench with the final build (gpurun --gpus 8); N = 1 is pass ai, N = 2 / 4 pass aj
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/scale_al_gpus.txt 2>&1
timeout 85 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711 \
  bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/bench_n8_al.json 2> gpurun_out/bench_n8_al.err; echo "n8 rc=$?"
cut -c1-300 gpurun_out/bench_n8_al.json
