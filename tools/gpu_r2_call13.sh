#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_update_gpu.py -x -q > gpurun_out/r2c13_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 gpurun_out/r2c13_pytest.log
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q -k "b200_dpvo_class" > gpurun_out/r2c13_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -n 3 gpurun_out/r2c13_pytest_seq.log
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --out gpurun_out/r2c13_b200_fps.json > gpurun_out/r2c13_b200_fps.log 2>&1; cut -c1-260 gpurun_out/r2c13_b200_fps.json; echo
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --no-loop-closure --out gpurun_out/r2c13_b200_fps_nolc.json > gpurun_out/r2c13_b200_fps_nolc.log 2>&1; cut -c1-260 gpurun_out/r2c13_b200_fps_nolc.json; echo
