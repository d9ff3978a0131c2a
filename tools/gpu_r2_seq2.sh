#!/bin/bash
# sequence parity (reference class on both back ends, this repo's class) + fps with / without loop closure
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q > gpurun_out/seq2_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/seq2_pytest.log
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --out gpurun_out/r02_sequence_fps_b200_dpvo_class.json > gpurun_out/seq_fps.log 2>&1; cut -c1-200 gpurun_out/r02_sequence_fps_b200_dpvo_class.json; echo
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --no-loop-closure --out gpurun_out/r02_sequence_fps_b200_dpvo_class_no_loop_closure.json > gpurun_out/seq_fps_nolc.log 2>&1; cut -c1-200 gpurun_out/r02_sequence_fps_b200_dpvo_class_no_loop_closure.json; echo
