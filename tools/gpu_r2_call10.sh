#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ba_gpu.py tests/test_encoder_gpu.py tests/test_update_gpu.py tests/test_fullframe_gpu.py -q > gpurun_out/r2c10_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r2c10_pytest.log
timeout 600 python tools/time_global_ba.py --frames 1000 > gpurun_out/r2c10_global_ba_1000.json 2> gpurun_out/r2c10_global_ba.err; cat gpurun_out/r2c10_global_ba_1000.json
timeout 600 python tools/time_global_ba.py --frames 300 --M 96 --loops 60 > gpurun_out/r2c10_global_ba_300.json 2>> gpurun_out/r2c10_global_ba.err; cat gpurun_out/r2c10_global_ba_300.json
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q -k "b200_dpvo_class" > gpurun_out/r2c10_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -n 5 gpurun_out/r2c10_pytest_seq.log
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --out gpurun_out/r2c10_b200_fps.json > gpurun_out/r2c10_b200_fps.log 2>&1; echo "fps rc=$?"; tail -n 2 gpurun_out/r2c10_b200_fps.log | cut -c1-400
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode profile --frames 1000 --out gpurun_out/r2c10_b200_profile.json > gpurun_out/r2c10_b200_profile.log 2>&1; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2c10_b200_profile.json'))
print(d["total_s"]); 
for k,v in d["stages"].items(): print("   %-34s calls %5d  s %7.3f  ms/call %8.3f"%(k,v["calls"],v["s"],v["ms_per_call"]))
PY
