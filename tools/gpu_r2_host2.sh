#!/bin/bash
# after the host-side trims: new unit tests, sequence parity, fps with / without loop closure, host profile
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pops_gpu.py tests/test_teacher_gpu.py tests/test_update_gpu.py tests/test_fullframe_gpu.py -x -q > gpurun_out/host2_unit.log 2>&1; echo "unit rc=$?"; tail -n 15 gpurun_out/host2_unit.log
bash tools/gpu_r2_seq2.sh
bash tools/gpu_r2_hostprof.sh
timeout 300 python tools/seq_gpu_busy.py --start 400 --count 100 --no-loop-closure --top 70 > gpurun_out/busy_nolc.log 2>&1
timeout 300 python tools/seq_gpu_busy.py --start 400 --count 100 --top 70 > gpurun_out/busy_lc.log 2>&1
