#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_reftrain_gpu.py -x -q > gpurun_out/r2c3_pytest_train.log 2>&1; echo "rc=$?" >> gpurun_out/r2c3_pytest_train.log
timeout 600 python tools/reftrain.py --backend dropin --no-amp --mode bench --steps 4 --warmup 2 --out gpurun_out/r2c3_train_dropin.json > gpurun_out/r2c3_train_dropin.log 2>&1
timeout 900 python tools/reftrain.py --backend ref --no-amp --mode bench --steps 3 --warmup 1 --out gpurun_out/r2c3_train_ref.json > gpurun_out/r2c3_train_ref.log 2>&1
tail -n 5 gpurun_out/r2c3_pytest_train.log gpurun_out/r2c3_pytest_ref.log gpurun_out/r2c3_train_dropin.log gpurun_out/r2c3_train_ref.log
timeout 300 python tools/time_corr_bwd.py --ref > gpurun_out/r2c3_corr_bwd_half.jsonl 2> gpurun_out/r2c3_corr_bwd.err
timeout 300 python tools/time_corr_bwd.py --ref --dtype float > gpurun_out/r2c3_corr_bwd_float.jsonl 2>> gpurun_out/r2c3_corr_bwd.err
timeout 600 python -m pytest tests/test_corr_gpu.py -x -q > gpurun_out/r2c3_pytest_corr.log 2>&1; echo "rc=$?" >> gpurun_out/r2c3_pytest_corr.log
tail -n 3 gpurun_out/r2c3_pytest_corr.log; cat gpurun_out/r2c3_corr_bwd_*.jsonl
