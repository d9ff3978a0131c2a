echo "row-major [45312, 384]:"; timeout 300 python tools/tma_probe.py 2>&1 | tail -8
echo "chunk-major (contiguous 16 KB boxes):"; timeout 300 python tools/tma_probe.py --chunked 2>&1 | tail -8
