mkdir -p gpurun_out
timeout 300 python tools/gemm_check.py > gpurun_out/gemm_check.log 2>&1; echo "rc=$?"; grep -c "max err" gpurun_out/gemm_check.log; grep "BAD\|timing\|ALL OK\|rror\|stream-l" gpurun_out/gemm_check.log | head -20
timeout 900 python -m pytest tests/test_update_gpu.py tests/test_ba_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/time_update.py > gpurun_out/time_update.json 2>gpurun_out/time_update.err; echo rc=$?; cat gpurun_out/time_update.json; tail -5 gpurun_out/time_update.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/update_launches.csv python tools/time_update.py --profile > gpurun_out/ncu_update.log 2>&1; echo ncu rc=$?
