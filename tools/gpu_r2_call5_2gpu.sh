#!/bin/bash
# gpurun --gpus 2: data-parallel training step (NCCL all-reduce on the data path) + the bench under torchrun
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${NGPU:-2}
NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
   tools/reftrain.py --backend dropin --no-amp --mode bench --steps 4 --warmup 2 --out gpurun_out/r2c5_train_${N}gpu.json > gpurun_out/r2c5_train_${N}gpu.log 2>&1
echo "train rc=$?"; grep -c "NCCL INFO" gpurun_out/r2c5_train_${N}gpu.log; grep -m3 "NVLS\|Connected all rings\|Channel 00" gpurun_out/r2c5_train_${N}gpu.log | cut -c1-200
cat gpurun_out/r2c5_train_${N}gpu.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 \
   bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2c5_bench_${N}gpu.json 2> gpurun_out/r2c5_bench_${N}gpu.err
echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c5_bench_${N}gpu.json')); print(d['n_gpus'], d['value'], d['e2e'], d['sustained'])"
timeout 600 python tools/reftrain.py --backend dropin --no-amp --mode bench --steps 4 --warmup 2 --out gpurun_out/r2c5_train_1gpu.json > gpurun_out/r2c5_train_1gpu.log 2>&1; cat gpurun_out/r2c5_train_1gpu.json
