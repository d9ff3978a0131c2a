#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_update_gpu.py tests/test_corr_gpu.py tests/test_fullframe_gpu.py -x -q > gpurun_out/r2c12_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2c12_pytest.log
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q -k "b200_dpvo_class" > gpurun_out/r2c12_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -n 3 gpurun_out/r2c12_pytest_seq.log
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --out gpurun_out/r2c12_b200_fps.json > gpurun_out/r2c12_b200_fps.log 2>&1; cut -c1-260 gpurun_out/r2c12_b200_fps.json; echo
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --no-loop-closure --out gpurun_out/r2c12_b200_fps_nolc.json > gpurun_out/r2c12_b200_fps_nolc.log 2>&1; cut -c1-260 gpurun_out/r2c12_b200_fps_nolc.json; echo
timeout 600 python bench.py --workload train --steps 3 --warmup 1 > gpurun_out/r2c12_bench_train.json 2> gpurun_out/r2c12_bench_train.err; echo "bench train rc=$?"; cut -c1-300 gpurun_out/r2c12_bench_train.json
