mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_corr_gpu.py tests/test_pops_gpu.py -m gpu -x -q 2>&1 | tail -4
for v in 0; do
  if [ $v = 1 ]; then export DPVO_CORR_NARROW=1; fi
  timeout 900 python bench.py --no-cpu-baseline --no-update-iteration > gpurun_out/bench_w$v.json 2>gpurun_out/bench_w$v.err; echo rc=$?
  python -c "
import json; d=json.load(open('gpurun_out/bench_w$v.json')); print('narrow=$v', d['value'], d['ms_per_step'], d['e2e']['value'], d['kernels_per_step'], d['roofline'])"
done
