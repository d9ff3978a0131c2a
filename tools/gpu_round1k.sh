mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_corr_gpu.py tests/test_ba_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_k.json 2>gpurun_out/bench_k.err; tail -c 1800 gpurun_out/bench_k.json
