#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/time_corr.py --ref > gpurun_out/r2c7_time_corr.jsonl 2> gpurun_out/r2c7_time_corr.err; cat gpurun_out/r2c7_time_corr.jsonl; tail -3 gpurun_out/r2c7_time_corr.err
timeout 300 python tools/time_corr_bwd.py > gpurun_out/r2c7_corr_bwd.jsonl 2> gpurun_out/r2c7_corr_bwd.err; cat gpurun_out/r2c7_corr_bwd.jsonl
timeout 300 python tools/time_corr_bwd.py --dtype float >> gpurun_out/r2c7_corr_bwd.jsonl 2>> gpurun_out/r2c7_corr_bwd.err; tail -2 gpurun_out/r2c7_corr_bwd.jsonl
timeout 900 python -m pytest tests/test_corr_gpu.py tests/test_fullsize_vs_reference_gpu.py -x -q > gpurun_out/r2c7_pytest_corr.log 2>&1; echo "pytest corr rc=$?"; tail -n 3 gpurun_out/r2c7_pytest_corr.log
rm -f gpurun_out/ba_accuracy.jsonl
timeout 1500 python -m pytest tests/test_ba_gpu.py tests/test_lietorch_gpu.py tests/test_viewer_gpu.py -q > gpurun_out/r2c7_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 25 gpurun_out/r2c7_pytest.log
