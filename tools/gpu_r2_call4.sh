#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c4_bench.json 2> gpurun_out/r2c4_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2c4_bench_ref.json 2> gpurun_out/r2c4_bench_ref.err; echo "bench ref rc=$?"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2c4_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 4 gpurun_out/r2c4_pytest_gpu.log; tail -n 5 gpurun_out/r2c4_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2c4_bench.json'))
for k in ("value","ms_per_step","e2e","sustained","gpu_launches","kernels_per_step","clocks"): print(k, d.get(k))
print("hot", d["hot_path"]); print("roof", d["roofline"]); print("refext", d.get("reference_cuda_ext")); 
s=d.get("sequence",{}); print("seq", {k:v for k,v in s.items() if k!="arms"}); print("cpu", d.get("cpu_baseline"))
PY
cut -c1-1500 gpurun_out/r2c4_bench_ref.json
