for v in 0 1; do echo "== AMC $v"; DPVO_LINEAR_AMC=$v timeout 300 python tools/gemm_check.py 2>&1 | grep "timing\|BAD\|ALL\|rror" ; DPVO_LINEAR_AMC=$v timeout 300 python tools/time_update.py; done
