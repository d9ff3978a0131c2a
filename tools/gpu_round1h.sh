mkdir -p gpurun_out
for n in 65 129; do echo "--- $n"; DPVO_DEBUG_SYNC=1 timeout 120 python tools/chol_debug.py $n 2>&1 | tail -1 | cut -c1-200; done
echo "--- all"; timeout 300 python tools/chol_debug.py 17 64 65 128 192 234 384 1000 1794 5994 2>&1 | tail -10 | cut -c1-200
timeout 900 python -m pytest tests/test_ba_gpu.py -m gpu -q 2>&1 | tail -4 | cut -c1-200
timeout 300 python tools/time_ba.py --ref 2>&1 | head -4
timeout 900 python tools/time_global_ba.py --frames 1000 --M 96 --loops 200 --ref 2>&1 | tail -1
