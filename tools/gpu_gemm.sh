mkdir -p gpurun_out
timeout 300 python tools/gemm_check.py > gpurun_out/gemm_check.log 2>&1; echo "rc=$?"; grep -c "max err" gpurun_out/gemm_check.log; grep "BAD\|timing\|ALL OK\|rror" gpurun_out/gemm_check.log | head -20
timeout 600 python -m pytest tests/test_update_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/time_update.py > gpurun_out/time_update.json 2>gpurun_out/time_update.err; echo rc=$?; cat gpurun_out/time_update.json; tail -5 gpurun_out/time_update.err
