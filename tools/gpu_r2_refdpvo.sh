#!/bin/bash
# Round 2: the reference's own DPVO class on both back ends (parity test, then 1000-frame FPS legs).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/r2_gpu.txt
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q > gpurun_out/r2_refdpvo_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_refdpvo_pytest.log
timeout 600 python -m pytest tests/test_corr_gpu.py -x -q > gpurun_out/r2_corr_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_corr_pytest.log
F=${FRAMES:-1000}
for arm in "ref reference" "dropin reference" "dropin b200"; do
  set -- $arm
  timeout 900 python tools/refdpvo.py --backend $1 --update $2 --mode fps --frames $F \
     --out gpurun_out/r2_fps_$1_$2.json > gpurun_out/r2_fps_$1_$2.log 2>&1
  echo "rc=$?" >> gpurun_out/r2_fps_$1_$2.log
done
tail -3 gpurun_out/r2_refdpvo_pytest.log gpurun_out/r2_corr_pytest.log
cat gpurun_out/r2_fps_*.json
