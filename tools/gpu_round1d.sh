set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/time_corr_bwd.py --ref --dtype half > gpurun_out/time_corr_bwd.jsonl 2>&1
timeout 600 python tools/time_corr_bwd.py --ref --dtype float >> gpurun_out/time_corr_bwd.jsonl 2>&1; cat gpurun_out/time_corr_bwd.jsonl
: > gpurun_out/corr_sweep.jsonl
for E in 1024 2048 4096 8192 16384 32768; do timeout 300 python tools/time_corr.py --edges $E --ref >> gpurun_out/corr_sweep.jsonl 2>&1; done
cat gpurun_out/corr_sweep.jsonl
timeout 600 python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 --iters 1 > gpurun_out/plain_gba.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_gba.csv python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 --iters 1 > gpurun_out/ncu_gba.log 2>&1
tail -2 gpurun_out/ncu_gba.log
