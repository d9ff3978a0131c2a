mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_m.json 2>gpurun_out/bench_m.err; echo rc=$?; python -c "
import json; d=json.load(open('gpurun_out/bench_m.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['kernels_per_step'], d['update_iteration'])"; tail -3 gpurun_out/bench_m.err
