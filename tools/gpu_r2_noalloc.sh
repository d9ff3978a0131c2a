#!/bin/bash
# allocation-free Cholesky entry point: alone in a fresh process, then after the BA tests in one process
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -n 5
timeout 900 python -m pytest tests/test_ba_gpu.py tests/test_pgo_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -n 5
python -m pytest tests --collect-only -q -m gpu 2>/dev/null | grep "::" | tail -n 3
