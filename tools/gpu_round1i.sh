mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python tools/time_ba.py --ref > gpurun_out/time_ba.jsonl 2>&1; head -4 gpurun_out/time_ba.jsonl
timeout 300 python tools/ba_accuracy.py > gpurun_out/ba_accuracy.jsonl 2>&1
timeout 900 python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 --ref > gpurun_out/global_ba_1000.json 2>&1; tail -1 gpurun_out/global_ba_1000.json
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cut -c1-900 gpurun_out/bench.json; tail -2 gpurun_out/bench.err
