timeout 1200 python -m pytest tests/test_ba_gpu.py -m gpu -x -q 2>&1 | tail -5
