set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_ba_gpu.py -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/time_ba.py --ref > gpurun_out/time_ba.jsonl 2>&1; head -4 gpurun_out/time_ba.jsonl
timeout 900 python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 --ref > gpurun_out/global_ba_1000.json 2>&1; tail -1 gpurun_out/global_ba_1000.json
timeout 300 python tools/ba_accuracy.py > gpurun_out/ba_accuracy.jsonl 2>&1
timeout 300 python tools/run_ba_once.py --frames 70 --window 64 > gpurun_out/plain_w64.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/launches_w64.csv python tools/run_ba_once.py --frames 70 --window 64 > gpurun_out/ncu_w64.log 2>&1
