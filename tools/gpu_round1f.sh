mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ba_gpu.py -m gpu -q 2>&1 | tail -3 | cut -c1-200
timeout 300 python tools/time_ba.py --ref 2>&1 | head -4
timeout 900 python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 2>&1 | tail -1
timeout 300 python tools/run_ba_once.py --frames 70 --window 64 > gpurun_out/plain_w64.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/launches_w64.csv python tools/run_ba_once.py --frames 70 --window 64 > gpurun_out/ncu_w64.log 2>&1
