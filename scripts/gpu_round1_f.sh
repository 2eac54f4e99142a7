#!/bin/bash
# sixth GPU pass (2 GPUs): N=1 and N=2 bench back to back, reference arm under torchrun
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/smi_f.txt
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 > gpurun_out/bench_n1_f.json 2> gpurun_out/bench_n1_f.err; echo "rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_n2_f.json 2> gpurun_out/bench_n2_f.err; echo "rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 3 > gpurun_out/bench_ref_n2_f.json 2> gpurun_out/bench_ref_n2_f.err; echo "rc=$?"
cat gpurun_out/bench_n1_f.json gpurun_out/bench_n2_f.json gpurun_out/bench_ref_n2_f.json | cut -c1-400
tail -3 gpurun_out/bench_n2_f.err
