mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_encoder_gpu.py -m gpu -x -q -s 2>&1 | tail -25
timeout 600 python tools/time_encoder.py 2>&1 | tail -3 | tee gpurun_out/time_encoder.json
