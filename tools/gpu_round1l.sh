mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 900 python bench.py > gpurun_out/bench_l.json 2>gpurun_out/bench_l.err; echo rc=$?; tail -c 2600 gpurun_out/bench_l.json; tail -3 gpurun_out/bench_l.err
