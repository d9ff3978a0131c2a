#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_corr_gpu.py tests/test_update_gpu.py tests/test_encoder_gpu.py -x -q > gpurun_out/r2c2_pytest_a.log 2>&1; echo "rc=$?" >> gpurun_out/r2c2_pytest_a.log
timeout 300 python tools/time_corr_bwd.py --ref > gpurun_out/r2c2_corr_bwd.jsonl 2> gpurun_out/r2c2_corr_bwd.err
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q > gpurun_out/r2c2_pytest_ref.log 2>&1; echo "rc=$?" >> gpurun_out/r2c2_pytest_ref.log
timeout 600 python tools/refdpvo.py --backend dropin --mode profile --frames 400 --out gpurun_out/r2c2_profile_dropin.json > gpurun_out/r2c2_profile_dropin.log 2>&1
timeout 600 python tools/refdpvo.py --backend dropin --update b200 --patchify b200 --mode profile --frames 400 --out gpurun_out/r2c2_profile_dropin_b200.json > gpurun_out/r2c2_profile_dropin_b200.log 2>&1
timeout 600 python tools/refdpvo.py --backend ref --mode profile --frames 400 --out gpurun_out/r2c2_profile_ref.json > gpurun_out/r2c2_profile_ref.log 2>&1
timeout 600 python tools/refdpvo.py --backend dropin --update b200 --mode fps --frames 1000 --out gpurun_out/r2c2_fps_dropin_b200upd.json > gpurun_out/r2c2_fps_dropin_b200upd.log 2>&1
timeout 600 python tools/refdpvo.py --backend dropin --update b200 --patchify b200 --mode fps --frames 1000 --out gpurun_out/r2c2_fps_dropin_b200all.json > gpurun_out/r2c2_fps_dropin_b200all.log 2>&1
tail -n 3 gpurun_out/r2c2_pytest_a.log gpurun_out/r2c2_pytest_ref.log
cat gpurun_out/r2c2_corr_bwd.jsonl gpurun_out/r2c2_fps_*.json
