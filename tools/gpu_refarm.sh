timeout 900 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -c 900
