#!/bin/bash
# config 5 at size after the session's work + the last global BA of the sequence, both bulk arithmetics
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python tools/time_global_ba.py --frames 1000 --ref 2>gpurun_out/gba1000.err | tee gpurun_out/r02_global_ba_1000.json | cut -c1-600
timeout 600 python tools/time_global_ba.py --frames 1500 2>gpurun_out/gba1500.err | tee gpurun_out/r02_global_ba_1500.json | cut -c1-400
timeout 600 python tools/trace_global_ba.py --sequence 1000 --tag final 2> gpurun_out/gba_trace.err | cut -c1-300
timeout 600 python tools/trace_global_ba.py --synthetic 1000 --tag final 2> gpurun_out/gba_trace.err | cut -c1-300
timeout 300 python tools/time_chol.py 2>&1 | cut -c1-200
