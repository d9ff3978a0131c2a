#!/bin/bash
# blocked Cholesky: tests first, then timing alone and inside the global BA
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ba_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -n 6
timeout 300 python tools/time_chol.py --fp32 2>&1 | cut -c1-1300
timeout 300 python tools/time_chol.py 2>&1 | cut -c1-1300
timeout 600 python tools/trace_global_ba.py --sequence 1000 ${GBA_TAG:+--tag $GBA_TAG} 2> gpurun_out/gba_trace.err | cut -c1-1500
