for cl in 1 8; do echo "== cluster $cl"; DPVO_LINEAR_CLUSTER=$cl timeout 300 python tools/gemm_check.py 2>&1 | grep "timing\|BAD\|ALL\|rror" ; done
