mkdir -p gpurun_out
DPVO_LINEAR_PROFILE=1 timeout 300 python tools/gemm_check.py --profile 2>&1 | grep "timing\|profile"
DPVO_LINEAR_PROFILE=1 timeout 300 python tools/gemm_check.py 2>&1 | grep "timing\|profile\|BAD\|ALL"
