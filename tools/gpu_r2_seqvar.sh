#!/bin/bash
# why does the sequence leg inside bench.py read lower than tools/refdpvo.py run alone?  same box, four ways
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 2>/dev/null | tail -n 1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['fps_loop'],1), round(d['fps'],1))"; }
echo "alone:            $(run)"
echo "alone again:      $(run)"
python - <<'PY'
import os, subprocess, sys, json
sys.path.insert(0, os.getcwd())
import bench
print("affinity before:", len(os.sched_getaffinity(0)), "cpus; numa-local list:", bench.bind_near_gpu(0), "->", len(os.sched_getaffinity(0)))
def run():
    r = subprocess.run([sys.executable, "tools/refdpvo.py", "--backend", "dropin", "--pipeline", "b200", "--mode", "fps", "--frames", "1000"], capture_output=True, text=True)
    d = json.loads(r.stdout.strip().splitlines()[-1]); return round(d["fps_loop"], 1), round(d["fps"], 1)
print("bound to the GPU's NUMA node:", run())
import torch
x = torch.zeros(1 << 28, device="cuda"); torch.cuda.synchronize()
print("bound + parent holds a CUDA context:", run())
PY
nproc; lscpu | grep -i "model name\|numa\|socket" | head -8
