timeout 600 python -m pytest tests/test_corr_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python bench.py --no-cpu-baseline --no-update-iteration 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['stage_ms'])"
