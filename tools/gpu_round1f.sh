mkdir -p gpurun_out
timeout 300 python tools/run_ba_once.py --frames 70 --window 64 > gpurun_out/plain_w64.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/launches_w64.csv python tools/run_ba_once.py --frames 70 --window 64 > gpurun_out/ncu_w64.log 2>&1
timeout 600 python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 --iters 1 > gpurun_out/plain_gba.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_gba2.csv python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 --iters 1 > gpurun_out/ncu_gba.log 2>&1
tail -1 gpurun_out/ncu_gba.log | cut -c1-200
