#!/bin/bash
# gpurun --gpus 8: data-parallel training step + the bench under torchrun on all 8 GPUs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${NGPU:-8}
NCCL_DEBUG=INFO timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 \
   bench.py --gpus $N --workload train --steps 3 --warmup 2 > gpurun_out/r2_train_${N}gpu.json 2> gpurun_out/r2_train_${N}gpu.log
echo "train rc=$?"; grep -m2 "NVLS\|Connected all" gpurun_out/r2_train_${N}gpu.log | cut -c1-160; tail -n 1 gpurun_out/r2_train_${N}gpu.json | cut -c1-500
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29552 \
   bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err
echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_${N}gpu.json')); print(d['n_gpus'], d['value'], d['e2e']['value'], d['sustained']['value'])"
