#!/bin/bash
# global BA: where the time goes (torch.profiler kernel records) + occupancy of S, on the sequence's own problem and the synthetic one
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for spec in "--sequence 1000" "--sequence 500" "--synthetic 1000"; do
  timeout 600 python tools/trace_global_ba.py $spec ${GBA_TAG:+--tag $GBA_TAG} 2> gpurun_out/gba_trace.err | cut -c1-1800
  echo "rc=$?"; tail -n 3 gpurun_out/gba_trace.err
done
