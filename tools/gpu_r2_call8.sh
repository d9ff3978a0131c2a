#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode parity --frames 120 --out gpurun_out/r2c8_b200_parity.npz > gpurun_out/r2c8_b200_parity.log 2>&1; echo "parity run rc=$?"; tail -n 12 gpurun_out/r2c8_b200_parity.log
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q -k "b200_dpvo_class" > gpurun_out/r2c8_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 gpurun_out/r2c8_pytest.log
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode profile --frames 400 --out gpurun_out/r2c8_b200_profile.json > gpurun_out/r2c8_b200_profile.log 2>&1; echo "profile rc=$?"; tail -n 3 gpurun_out/r2c8_b200_profile.log | cut -c1-1500
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --out gpurun_out/r2c8_b200_fps.json > gpurun_out/r2c8_b200_fps.log 2>&1; echo "fps rc=$?"; tail -n 3 gpurun_out/r2c8_b200_fps.log | cut -c1-800
