#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python tools/seq_gpu_busy.py --start 400 --count 200 --host > gpurun_out/hostprof_lc.log 2>&1; tail -n 2 gpurun_out/hostprof_lc.log
timeout 600 python tools/seq_gpu_busy.py --start 400 --count 200 --host --no-loop-closure > gpurun_out/hostprof_nolc.log 2>&1; tail -n 2 gpurun_out/hostprof_nolc.log
