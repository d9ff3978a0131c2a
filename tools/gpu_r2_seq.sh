#!/bin/bash
# the 1,000-frame sequence on this repo's DPVO class: stage profile, fps with and without loop closure
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ba_gpu.py -x -q -m gpu -k "degenerate or eff_impl" 2>&1 | tail -n 3
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode profile --frames 1000 --out gpurun_out/r02_sequence_profile_b200_dpvo_class.json > gpurun_out/seq_profile.log 2>&1; echo "profile rc=$?"
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --out gpurun_out/r02_sequence_fps_b200_dpvo_class.json > gpurun_out/seq_fps.log 2>&1; cut -c1-200 gpurun_out/r02_sequence_fps_b200_dpvo_class.json; echo
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --no-loop-closure --out gpurun_out/r02_sequence_fps_b200_dpvo_class_no_loop_closure.json > gpurun_out/seq_fps_nolc.log 2>&1; cut -c1-200 gpurun_out/r02_sequence_fps_b200_dpvo_class_no_loop_closure.json; echo
