mkdir -p gpurun_out
for v in b n b n b n; do
  if [ $v = n ]; then export DPVO_BENCH_NO_BIND=1; else unset DPVO_BENCH_NO_BIND; fi
  timeout 600 python bench.py --no-cpu-baseline --no-update-iteration > gpurun_out/bench_ab.json 2>gpurun_out/bench_ab.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_ab.json')); print('$v', round(d['value']), round(d['e2e']['value']), d['config']['host_cpus'], round(d['e2e']['h2d_gbps_if_copy_bound'],1))"
done
nproc; cat /sys/devices/system/node/online 2>/dev/null; python -c "import os; print(len(os.sched_getaffinity(0)))"
