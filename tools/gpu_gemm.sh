mkdir -p gpurun_out
timeout 300 python tools/gemm_check.py > gpurun_out/gemm_check.log 2>&1; echo "rc=$?"; tail -60 gpurun_out/gemm_check.log
