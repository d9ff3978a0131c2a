mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_edges_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/time_edges.py 2>&1 | tail -3 | tee gpurun_out/time_edges.json
