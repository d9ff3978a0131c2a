mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_ba_gpu.py tests/test_update_gpu.py -m gpu -x -q 2>&1 | tail -4
for v in 1 0 1 0; do
  DPVO_BA_SOLVE_WARP=$v timeout 600 python bench.py --no-cpu-baseline --no-update-iteration > gpurun_out/bench_s$v.json 2>gpurun_out/bench_s$v.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_s$v.json')); print('warp=$v', round(d['value'],1), d['ms_per_step'], d['stage_ms']['ba'], round(d['e2e']['value']))"
done
