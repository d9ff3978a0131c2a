#!/bin/bash
# one-read keyframe(): unit test, sequence parity (with verify_plan on our class), fps with loop closure
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pops_gpu.py -x -q > gpurun_out/plan_unit.log 2>&1; echo "unit rc=$?"; tail -n 12 gpurun_out/plan_unit.log
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q > gpurun_out/seq2_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/seq2_pytest.log
timeout 300 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 2>/dev/null | tail -n 1 | cut -c1-120
timeout 300 python tools/refdpvo.py --backend dropin --pipeline b200 --mode parity --frames 400 2>&1 | tail -n 2 | cut -c1-200
